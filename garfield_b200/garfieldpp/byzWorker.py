"""`ByzWorker` — a worker that replies with an attack instead of its honest gradient.

API parity with /root/reference/pytorch_impl/libs/garfieldpp/byzWorker.py:46-143 (attacks
random, reverse, drop, lie, empire; `fw` colluders). Attack vectors are generated on the
worker's device by `garfield_b200.ops.attacks` (Philox kernels), not with CPU tensor ops.
`strict_reference=True` reproduces two reference quirks: `drop` is a no-op (Q5) and `lie` with a
single colluder yields NaN (unbiased std of one sample).
"""

from __future__ import annotations

from .worker import Worker
from ..ops import attacks

__all__ = ["ByzWorker"]


class ByzWorker(Worker):
    def __init__(self, rank, world_size, num_workers, batch_size, model, dataset, loss, attack, fw=1, train_size=None,
                 device=None, strict_reference=False, z=1.035, eps=10.0, drop_p=0.3, reverse_coeff=100.0):
        super().__init__(rank, world_size, num_workers, batch_size, model, dataset, loss, train_size, device=device)
        self.fw = fw
        self.strict_reference = strict_reference
        self.z, self.eps, self.drop_p, self.reverse_coeff = z, eps, drop_p, reverse_coeff
        table = {"random": self.random_attack, "reverse": self.reverse_attack, "drop": self.partial_drop_attack,
                 "lie": self.little_is_enough_attack, "empire": self.fall_empires_attack}
        assert attack in table, "The requested attack is not implemented; available attacks are:" + str(table.keys())
        self.attack_name = attack
        self.attack = table[attack]

    def compute_gradients(self, iter_num, model, out=None):
        rank, grad, loss = self.attack(iter_num, model)
        if out is not None:
            out.copy_(grad)
            grad = out
        return rank, grad, loss

    def _honest(self, iter_num, model):
        rank, grad, loss = super().compute_gradients(iter_num, model)
        return rank, grad.clone(), loss

    def random_attack(self, iter_num, model):
        rank, grad, loss = self._honest(iter_num, model)
        return rank, attacks.random_like(grad), loss

    def reverse_attack(self, iter_num, model):
        rank, grad, loss = self._honest(iter_num, model)
        return rank, attacks.reverse(grad, self.reverse_coeff), loss

    def partial_drop_attack(self, iter_num, model):
        rank, grad, loss = self._honest(iter_num, model)
        return rank, attacks.drop(grad, self.drop_p, strict_reference=self.strict_reference), loss

    def _estimates(self, iter_num, model):
        """Honest gradients on batches iter, iter+1, … simulating `fw` colluding workers."""
        rank, grad, loss = self._honest(iter_num, model)
        estimates = [self._honest(iter_num + i + 1, model)[1] for i in range(self.fw - 1)]
        estimates.append(grad)
        return rank, estimates, loss

    def little_is_enough_attack(self, iter_num, model):
        rank, estimates, loss = self._estimates(iter_num, model)
        return rank, attacks.lie(estimates, self.z, strict_reference=self.strict_reference), loss

    def fall_empires_attack(self, iter_num, model):
        rank, estimates, loss = self._estimates(iter_num, model)
        return rank, attacks.empire(estimates, self.eps), loss
