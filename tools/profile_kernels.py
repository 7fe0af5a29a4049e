#!/usr/bin/env python
"""One launch of every kernel of the library at BASELINE sizes, for ncu:
    ncu --set full --clock-control none --import-source on -k regex:ps_ -o out python tools/profile_kernels.py
Prints the order of launches so the report rows can be labelled."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402
from kubeflow_controller_b200 import sharded  # noqa: E402

DEV = "cuda:0"
N = 100_000_000


def main():
    ctx = pkg.Context(1)
    var = torch.randn(N, device=DEV)
    m = torch.zeros(N, device=DEV)
    v = torch.full((N,), 1e-4, device=DEV)
    g32 = [torch.randn(N, device=DEV) * 0.01 for _ in range(8)]
    hp = pkg.HParams(lr=0.01, momentum=0.9)
    order = []
    ctx.apply_sgd(var, g32, hp); order.append("ps_apply_ldg SGD f32 N=8 (AUTO: 256-bit)")
    ctx.apply_momentum(var, m, g32, hp); order.append("ps_apply_ldg Momentum f32 N=8")
    ctx.apply_adam(var, m, v, g32, hp); order.append("ps_apply_ldg Adam f32 N=8")
    out = torch.zeros(1, device=DEV)
    ctx.grad_sumsq(g32, out); order.append("ps_sumsq f32 N=8")
    ctx.set_kernel(pkg.KERNEL_TMA)
    ctx.apply_sgd(var, g32, hp); order.append("ps_apply_tma SGD f32 N=8")
    ctx.set_kernel(pkg.KERNEL_AUTO)
    g16 = [g.to(torch.bfloat16) for g in g32[:1]]
    ctx.apply_adam(var[:41_892_736], m[:41_892_736], v[:41_892_736],
                   [g16[0][:41_892_736]], hp)
    order.append("ps_apply_ldg Adam bf16 N=1 (config 4 shard, 41.9 M)")
    torch.cuda.synchronize()
    del m, v, g16
    # the sharded kernels, two ranks emulated on this GPU (peer stores are local here)
    ranks = sharded.local_group(2, N, pkg.OPT_SGD, lambda: pkg.HParams(lr=0.01))
    grads = [g32, g32]
    eps = [x.push(g) for x, g in zip(ranks, grads)]
    order += ["ps_push f32 N=8 rank 0 of 2", "ps_push f32 N=8 rank 1 of 2"]
    for x in ranks:
        x.apply_pull(16, epoch=eps[0])
    order += ["ps_apply_pull SGD rank 0 of 2", "ps_apply_pull SGD rank 1 of 2"]
    for x in ranks:
        x.wait_pull(epoch=eps[0])
    order += ["ps_wait_pull", "ps_wait_pull"]
    sharded.local_group_step_fused(ranks, grads, n_slices=4)
    order.append("ps_step_fused SGD f32 N=8, 2 ranks in one cooperative launch, 4 slices")
    sharded.local_group_step_fused(ranks, grads, n_slices=8, overlap=pkg.accel.OVERLAP_SWITCH)
    order.append("ps_step_switch SGD f32 N=8, 2 ranks in one cooperative launch, 8 slices "
                 "(peer-load form: no NVSwitch inside one GPU)")
    torch.cuda.synchronize()
    del ranks, grads, g32
    torch.cuda.empty_cache()
    # config 4's push with raw-bf16 stage slots: one bf16 gradient per rank, 8 ranks on this GPU
    n4 = 41_892_736 * 8 // 4        # a quarter of BERT-large keeps 8 emulated ranks in memory
    ranks = sharded.local_group(8, n4, pkg.OPT_ADAM, lambda: pkg.HParams(lr=1e-4))
    g16 = [[(torch.randn(ranks[0].padded, device=DEV) * 0.01).to(torch.bfloat16)] for _ in ranks]
    sharded.local_group_step_fused(ranks, g16, n_slices=2)
    order.append("ps_step_fused Adam bf16 N=1 raw-bf16 stage, 8 ranks in one cooperative launch")
    torch.cuda.synchronize()
    for i, o in enumerate(order):
        print(i, o)


if __name__ == "__main__":
    main()
