"""omegaconf stand-in: the names the reference touches at import time / in type annotations (config/__init__.py:6,
scene/mlp_scene.py:6).  The hydra CLI layer is out of scope; configs are constructed as dataclasses directly."""


class DictConfig(dict):
  def __getattr__(self, k):
    try:
      return self[k]
    except KeyError as e:
      raise AttributeError(k) from e


class ListConfig(list):
  pass


class OmegaConf:
  _resolvers = {}

  @classmethod
  def register_new_resolver(cls, name, fn, replace=False, **kwargs):
    cls._resolvers[name] = fn

  @staticmethod
  def create(obj=None):
    return DictConfig(obj or {})

  @staticmethod
  def to_container(cfg, resolve=True, **kwargs):
    return dict(cfg) if isinstance(cfg, dict) else list(cfg)

  @staticmethod
  def to_yaml(cfg, **kwargs):
    import yaml
    return yaml.safe_dump(OmegaConf.to_container(cfg))

  @staticmethod
  def is_config(obj):
    return isinstance(obj, (DictConfig, ListConfig))
