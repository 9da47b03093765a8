"""`ByzServer` — a server replica that lies about its model when peers pull it.

API parity with /root/reference/pytorch_impl/libs/garfieldpp/byzServer.py:53-107 (attacks
random, reverse, drop). The reference version is internally inconsistent (it treats the
`(rank, tensor)` pair as a tensor); here `get_model()` always returns
`(rank, flat)` and the attack is applied to `flat` on the device.
"""

from __future__ import annotations

from .server import Server
from ..ops import attacks

__all__ = ["ByzServer"]


class ByzServer(Server):
    def __init__(self, rank, world_size, num_workers, num_ps, byz_wrk, byz_ps, wrk_base_name, ps_base_name, batch, model,
                 dataset, optimizer, attack, *args, strict_reference=False, **kwargs):
        super().__init__(rank, world_size, num_workers, num_ps, byz_wrk, byz_ps, wrk_base_name, ps_base_name, batch, model,
                         dataset, optimizer, *args, **kwargs)
        table = {"random": self.random_attack, "reverse": self.reverse_attack, "drop": self.partial_drop_attack}
        assert attack in table, "The requested attack is not implemeneted; available attacks are:" + str(table.keys())
        self.attack_name = attack
        self.attack = table[attack]
        self.strict_reference = strict_reference

    def get_model(self):
        return self.attack()

    def random_attack(self):
        rank, flat = super().get_model()
        return rank, attacks.random_like(flat)

    def reverse_attack(self):
        rank, flat = super().get_model()
        return rank, attacks.reverse(flat, 100.0)

    def partial_drop_attack(self):
        rank, flat = super().get_model()
        return rank, attacks.drop(flat, 0.3, strict_reference=self.strict_reference)
