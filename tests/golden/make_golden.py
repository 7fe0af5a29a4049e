"""Generate tests/golden/ps_step_golden.npz from the CPU oracle.

The reference's arithmetic (TensorFlow 1.4.0) cannot be imported or run in this
image, so these fixtures are produced by oracle/ps_oracle.py — the restatement
pinned by the hand-derived known-answer vectors in kat_survey.json — and serve
as a regression pin for both the oracle and the CUDA path (the GPU tests replay
them through the C ABI). Run from the repo root: python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ps_oracle as O  # noqa: E402

F = np.float32
HP_KEYS = ["lr", "momentum", "beta1", "beta2", "epsilon", "beta1_power", "beta2_power",
           "clip_norm", "use_nesterov", "aggregate"]


def case(name, opt, n, workers, hp, bf16, seed, out):
    rng = np.random.default_rng(seed)
    var = rng.standard_normal(n).astype(F)
    s0 = (0.1 * rng.standard_normal(n)).astype(F)
    s1 = np.abs(0.01 * rng.standard_normal(n)).astype(F)
    grads = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(workers)]
    if bf16:
        grads = [O.f32_to_bf16_bits(g) for g in grads]
    out[f"{name}.opt"] = opt
    out[f"{name}.workers"] = workers
    out[f"{name}.hp_keys"] = np.array(HP_KEYS)
    out[f"{name}.hp_vals"] = np.array([float(getattr(hp, k)) for k in HP_KEYS])
    out[f"{name}.var"], out[f"{name}.s0"], out[f"{name}.s1"] = var.copy(), s0.copy(), s1.copy()
    for w, g in enumerate(grads):
        out[f"{name}.g{w}"] = g
    _, ss = O.numpy_ps_step(opt, var, s0, s1, grads, hp)
    out[f"{name}.sumsq"] = F(0.0 if ss is None else ss)
    out[f"{name}.var_out"], out[f"{name}.s0_out"], out[f"{name}.s1_out"] = var, s0, s1


def main():
    out = {}
    case("sgd", O.SGD, 2051, 8, O.HParams(lr=0.01), False, 1, out)
    case("momentum", O.MOMENTUM, 1537, 3, O.HParams(lr=0.1, momentum=0.9, use_nesterov=True),
         False, 2, out)
    hp = O.HParams(lr=1e-4, aggregate=O.AGG_MEAN, clip_norm=0.01)
    for _ in range(999):   # step t = 1000 of BASELINE config 4
        hp.advance()
    case("adam_bf16_mean_clip", O.ADAM, 3079, 5, hp, True, 3, out)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ps_step_golden.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
