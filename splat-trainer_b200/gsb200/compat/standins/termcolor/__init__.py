"""termcolor stand-in (reference config/__init__.py:8, trainer/trainer.py:24): no colours."""


def colored(text, *args, **kwargs):
  return str(text)


def cprint(text, *args, **kwargs):
  print(text)
