"""numpy attack library — parity with tensorflow_impl/libs/attacker.py:35-128 (`Attacker(attack)`
with Random, Reverse, PartialDrop, LittleIsEnough, FallEmpires; note the TF side uses the BIASED
standard deviation for "little is enough")."""

from __future__ import annotations

import numpy as np

__all__ = ["Attacker"]


class Attacker:
    def __init__(self, attack):
        table = {"Random": self.random_attack, "Reverse": self.reverse_attack, "PartialDrop": self.partial_drop_attack,
                 "LittleIsEnough": self.little_is_enough_attack, "FallEmpires": self.fall_empires_attack}
        assert attack in table, "unknown attack %r; available: %s" % (attack, sorted(table))
        self.attack_strategy = table[attack]

    def attack(self, **kwargs):
        return self.attack_strategy(**kwargs)

    @staticmethod
    def random_attack(gradient):
        return np.random.random(gradient.shape).astype(np.float32)

    @staticmethod
    def reverse_attack(gradient, coeff=100):
        return gradient * coeff * (-1.)

    @staticmethod
    def partial_drop_attack(gradient, probability):
        mask = np.random.random(gradient.shape) < probability
        return np.where(mask, 0, gradient).astype(gradient.dtype)

    @staticmethod
    def little_is_enough_attack(gradient, byz_gradients, z=1.035):
        stack = np.stack(list(byz_gradients) + [gradient])
        return (stack.mean(axis=0) + z * stack.std(axis=0)).astype(np.float32)

    @staticmethod
    def fall_empires_attack(gradient, byz_gradients, eps=10):
        stack = np.stack(list(byz_gradients) + [gradient])
        return (-eps * stack.mean(axis=0)).astype(np.float32)
