"""TaichiQueue (reference scripts/train_scan.py:237, scene/mlp_scene.py:417): upstream serialises taichi launches on
one worker thread.  libgsb200 is re-entrant and enqueues on the caller's CUDA stream, so calls run inline."""


class TaichiQueue:
  _initialised = False

  @classmethod
  def init(cls, arch=None, debug=False, device_memory_GB=None, threaded=False, **kwargs):
    from gsb200 import lib
    lib.load()  # fail here, loudly, if the CUDA library is missing
    cls._initialised = True

  @classmethod
  def run_sync(cls, f, *args, **kwargs):
    return f(*args, **kwargs)

  @classmethod
  def run_async(cls, f, *args, **kwargs):
    from concurrent.futures import Future
    fut = Future()
    try:
      fut.set_result(f(*args, **kwargs))
    except BaseException as e:  # noqa: BLE001
      fut.set_exception(e)
    return fut

  @classmethod
  def stop(cls):
    cls._initialised = False
