"""wonderwords stand-in (reference config/__init__.py:9,188-195: random run names)."""
import random


class RandomWord:
  _W = ("amber", "birch", "cedar", "delta", "ember", "fjord", "glade", "heron", "iris", "jade", "kelp", "lotus")

  def word(self, **kwargs):
    return random.choice(self._W)

  def random_words(self, amount=1, **kwargs):
    return [self.word() for _ in range(amount)]
