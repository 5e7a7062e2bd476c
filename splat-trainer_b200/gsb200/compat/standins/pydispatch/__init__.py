"""pydispatch stand-in (reference trainer/trainer.py:19: Dispatcher with emit / bind for progress events)."""


class Dispatcher:
  _events_ = []

  def bind(self, **handlers):
    table = self.__dict__.setdefault("_handlers", {})
    for name, fn in handlers.items():
      table.setdefault(name, []).append(fn)

  def unbind(self, *fns):
    for hs in self.__dict__.get("_handlers", {}).values():
      for fn in fns:
        if fn in hs:
          hs.remove(fn)

  def emit(self, name, *args, **kwargs):
    for fn in list(self.__dict__.get("_handlers", {}).get(name, [])):
      fn(*args, **kwargs)
