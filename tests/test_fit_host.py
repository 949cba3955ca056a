"""Host logic of b200cgan.fit on the CPU: the ReduceLROnPlateau restatement against torch's own scheduler (the reference
builds it with mode='max', factor=0.5, patience=5, cgan/cgan_train_enhanced.py:260-261) on metric sequences with plateaus,
improvements inside the relative threshold, and negative values (delta-IoU can be negative)."""
import random

import torch

from b200cgan.fit import ReduceLROnPlateau


class _Opt:
    def __init__(self, lr):
        self.lr = lr


def test_reduce_lr_on_plateau_matches_torch():
    rnd = random.Random(3)
    seqs = [[0.01] * 30, [0.001 * i for i in range(8)] + [0.0069] * 20 + [0.02] + [0.0] * 15,
            [rnd.uniform(-0.01, 0.02) for _ in range(60)], [-0.005 + 1e-7 * i for i in range(40)]]
    for seq in seqs:
        p = torch.nn.Parameter(torch.zeros(1))
        topt = torch.optim.Adam([p], lr=2e-4)
        tsch = torch.optim.lr_scheduler.ReduceLROnPlateau(topt, mode="max", factor=0.5, patience=5)
        mine_opt = _Opt(2e-4)
        mine = ReduceLROnPlateau(mine_opt, mode="max", factor=0.5, patience=5)
        for m in seq:
            tsch.step(m)
            mine.step(m)
            assert abs(topt.param_groups[0]["lr"] - mine_opt.lr) < 1e-15, (m, topt.param_groups[0]["lr"], mine_opt.lr)
        if seq is seqs[0]:
            assert mine_opt.lr < 2e-4  # a flat metric does reduce the rate
