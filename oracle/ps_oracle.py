"""CPU oracle of the parameter-server step — TEST INFRASTRUCTURE, NOT PRODUCT.

Two restatements of the same published TF-1.4 algorithms, which must agree bit
for bit with each other:

* ``numpy_*``  : numpy float32, one ufunc per operation (each operation rounds
  on its own, no FMA), one array pass per TF statement;
* ``COracle``  : ctypes face of oracle/cpu_ps.c (threaded; also the timed CPU
  baseline of bench.py).

PARITY UNPINNED: the arithmetic lives in TensorFlow v1.4.0 (image tag at
/root/reference/examples/tfjob/dist.yml:13,32), which is neither in
/root/reference nor importable here, and the reference holds no test or golden
vector for the path (only pkg/checker/checker_test.go:10-38). The formulas
follow tensorflow/core/kernels/training_ops.cc (ApplyGradientDescent,
ApplyMomentum, ApplyAdamNonCuda), conditional_accumulator.h, clip_ops.py and
adam.py of TF 1.4.0, anchored on the reference call sites
examples/workdir/mnist_replica.py:170,172-184,256 and mnist_softmax.py:59; the
known-answer vectors of SURVEY.md §8(c) are in tests/golden/kat_survey.json.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass

import numpy as np

F32 = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))

SGD, MOMENTUM, ADAM = 0, 1, 2
DT_F32, DT_BF16 = 0, 1
AGG_SUM, AGG_MEAN = 0, 1


@dataclass
class HParams:
    lr: float = 0.01
    momentum: float = 0.9
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8
    beta1_power: float = 0.9    # TF adam.py _create_slots: initial value beta1
    beta2_power: float = 0.999  # initial value beta2
    clip_norm: float = 0.0
    use_nesterov: bool = False
    aggregate: int = AGG_SUM

    def advance(self) -> None:
        """TF-1.4 AdamOptimizer._finish: beta_power *= beta, in fp32."""
        self.beta1_power = float(F32(self.beta1_power) * F32(self.beta1))
        self.beta2_power = float(F32(self.beta2_power) * F32(self.beta2))


# ---- bf16 helpers (inputs of BASELINE config 4) ------------------------------
def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16, returned as uint16 bit patterns."""
    u = np.ascontiguousarray(x, dtype=F32).view(np.uint32).astype(np.uint64)
    rounded = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return rounded.astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(F32)


# ---- numpy restatement -------------------------------------------------------
def numpy_aggregate(grads: list[np.ndarray], mean: bool) -> np.ndarray:
    """ConditionalAccumulator: acc = g0; acc += g_w in worker order; take_grad
    divides by the count (SyncReplicasOptimizer, mnist_replica.py:172-182)."""
    def as_f32(g):
        return bf16_bits_to_f32(g) if g.dtype == np.uint16 else g.astype(F32, copy=False)
    acc = as_f32(grads[0]).copy()
    for g in grads[1:]:
        acc = acc + as_f32(g)
    if mean:
        acc = acc / F32(len(grads))
    return acc


def numpy_sumsq(g: np.ndarray) -> np.float32:
    """sum(g^2) for global_norm. Reduction order is unspecified in TF (Eigen tree
    reduction), so this uses fp64 accumulation of fp32 products and callers
    compare with the path's tolerance."""
    sq = g * g
    return F32(np.sum(sq.astype(np.float64)))


def numpy_clip_scale(sumsq: np.float32, clip_norm: float) -> np.float32:
    """clip_by_global_norm: clip_norm * min(1/global_norm, 1/clip_norm)."""
    c = F32(clip_norm)
    gn = np.sqrt(F32(sumsq))
    with np.errstate(divide="ignore"):
        return c * np.minimum(F32(1.0) / gn, F32(1.0) / c)


def numpy_adam_alpha(hp: HParams) -> np.float32:
    num = F32(hp.lr) * np.sqrt(F32(1.0) - F32(hp.beta2_power))
    return num / (F32(1.0) - F32(hp.beta1_power))


def numpy_apply(optimizer: int, var: np.ndarray, slot0, slot1, g: np.ndarray, hp: HParams):
    """In-place TF-1.4 CPU apply; statement per pass."""
    lr = F32(hp.lr)
    if optimizer == SGD:
        var[...] = var - g * lr
    elif optimizer == MOMENTUM:
        mu = F32(hp.momentum)
        slot0[...] = slot0 * mu + g
        if hp.use_nesterov:
            var[...] = var - (g * lr + (slot0 * mu) * lr)
        else:
            var[...] = var - slot0 * lr
    elif optimizer == ADAM:
        b1 = F32(hp.beta1)
        omb1 = F32(1.0) - b1
        omb2 = F32(1.0) - F32(hp.beta2)
        alpha = numpy_adam_alpha(hp)
        eps = F32(hp.epsilon)
        slot0[...] = slot0 + (g - slot0) * omb1
        slot1[...] = slot1 + (g * g - slot1) * omb2
        if hp.use_nesterov:
            var[...] = var - ((g * omb1 + b1 * slot0) * alpha) / (np.sqrt(slot1) + eps)
        else:
            var[...] = var - (slot0 * alpha) / (np.sqrt(slot1) + eps)
    else:
        raise ValueError(f"unknown optimizer {optimizer}")


def numpy_ps_step(optimizer: int, var, slot0, slot1, grads, hp: HParams, sumsq=None):
    """Aggregate -> (clip) -> apply, in place. Returns (aggregated gradient as
    applied, sum of squares used for the clip or None)."""
    g = numpy_aggregate(grads, hp.aggregate == AGG_MEAN)
    ss = None
    if hp.clip_norm > 0:
        ss = numpy_sumsq(g) if sumsq is None else F32(sumsq)
        g = g * numpy_clip_scale(ss, hp.clip_norm)
    numpy_apply(optimizer, var, slot0, slot1, g, hp)
    return g, ss


# ---- C restatement -------------------------------------------------------------
class _CHParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in
                ("lr", "momentum", "beta1", "beta2", "epsilon", "beta1_power", "beta2_power",
                 "clip_norm")] + [("use_nesterov", ctypes.c_int32), ("aggregate", ctypes.c_int32)]


def _c_hparams(hp: HParams) -> _CHParams:
    return _CHParams(hp.lr, hp.momentum, hp.beta1, hp.beta2, hp.epsilon, hp.beta1_power,
                     hp.beta2_power, hp.clip_norm, int(hp.use_nesterov), int(hp.aggregate))


def build_c_oracle(force: bool = False) -> str:
    lib = os.path.join(HERE, "liboracle_ps.so")
    src = os.path.join(HERE, "cpu_ps.c")
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B" if force else "-s", "liboracle_ps.so"],
                       check=True, capture_output=True)
    return lib


def build_native_c_oracle() -> str | None:
    """The same source built -march=native ON THIS MACHINE (always rebuilt: a copy built on
    another CPU could hold instructions this one lacks). Used by bench.py's timed CPU baseline
    only; results are identical to the portable build (-ffp-contract=off in both)."""
    lib = os.path.join(HERE, "liboracle_ps_native.so")
    p = subprocess.run(["make", "-C", HERE, "-B", "liboracle_ps_native.so"], capture_output=True)
    return lib if p.returncode == 0 and os.path.exists(lib) else None


class COracle:
    def __init__(self, threads: int | None = None, native: bool = False):
        self.build = "x86-64-v3"
        path = build_native_c_oracle() if native else None
        if path:
            self.build = "native"
        else:
            path = build_c_oracle()
        try:
            self.lib = ctypes.CDLL(path)
        except OSError:
            self.build = "x86-64-v3"
            self.lib = ctypes.CDLL(build_c_oracle(force=True))
        self.threads = threads or (os.cpu_count() or 1)
        L = self.lib
        fp = ctypes.POINTER(ctypes.c_float)
        L.oracle_ps_step.restype = ctypes.c_float
        L.oracle_ps_step.argtypes = [ctypes.c_int, fp, fp, fp, ctypes.POINTER(ctypes.c_void_p),
                                     ctypes.c_int, ctypes.c_int, ctypes.c_uint64,
                                     ctypes.POINTER(_CHParams), fp, ctypes.c_int]
        L.oracle_sumsq.restype = ctypes.c_float
        L.oracle_sumsq.argtypes = [fp, ctypes.c_uint64, ctypes.c_int]
        L.oracle_aggregate.restype = None
        L.oracle_aggregate.argtypes = [fp, ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                       ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_int]

    @staticmethod
    def _fp(a):
        if a is None:
            return None
        assert a.dtype == F32 and a.flags.c_contiguous
        return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))

    def ps_step(self, optimizer, var, slot0, slot1, grads, hp: HParams, acc=None) -> float:
        """In place on var/slot0/slot1. grads: float32 arrays, or uint16 (bf16 bits)."""
        n = var.size
        dtype = DT_BF16 if grads[0].dtype == np.uint16 else DT_F32
        for g in grads:
            assert g.size == n and g.flags.c_contiguous
            assert g.dtype == (np.uint16 if dtype == DT_BF16 else F32)
        if acc is None:
            acc = np.empty(n, dtype=F32)
        ptrs = (ctypes.c_void_p * len(grads))(*[g.ctypes.data for g in grads])
        chp = _c_hparams(hp)
        return float(self.lib.oracle_ps_step(optimizer, self._fp(var), self._fp(slot0),
                                             self._fp(slot1), ptrs, len(grads), dtype, n,
                                             ctypes.byref(chp), self._fp(acc), self.threads))

    def aggregate(self, grads, mean: bool) -> np.ndarray:
        n = grads[0].size
        dtype = DT_BF16 if grads[0].dtype == np.uint16 else DT_F32
        acc = np.empty(n, dtype=F32)
        ptrs = (ctypes.c_void_p * len(grads))(*[g.ctypes.data for g in grads])
        self.lib.oracle_aggregate(self._fp(acc), ptrs, len(grads), dtype, n, int(mean),
                                  self.threads)
        return acc

    def sumsq(self, g) -> float:
        return float(self.lib.oracle_sumsq(self._fp(g), g.size, self.threads))
