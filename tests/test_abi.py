"""The C-ABI library loads, exports every symbol include/accel_ps.h declares, checks
its arguments, and fails loudly (never computes on the CPU) without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "accel_ps.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(accel_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_list_agree(pkg):
    assert _declared_symbols() == sorted(pkg.accel.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in _declared_symbols():
        assert hasattr(lib, name), f"libaccel_ps.so does not export {name}"


def test_version(pkg):
    assert pkg.accel.version() == 10200


def test_hparams_layout_matches_header(pkg):
    assert ctypes.sizeof(pkg.HParams) == 40
    hp = pkg.HParams(lr=0.5, beta1=0.8, beta2=0.99)
    assert hp.beta1_power == pytest.approx(0.8) and hp.beta2_power == pytest.approx(0.99)
    pkg.accel.adam_advance(hp)
    import numpy as np
    assert hp.beta1_power == float(np.float32(0.8) * np.float32(0.8))
    assert hp.beta2_power == float(np.float32(0.99) * np.float32(0.99))


def test_no_oracle_in_product(lib):
    """The product library must not link or reference the oracle."""
    path = os.path.join(ROOT, "kubeflow-controller_b200", "libaccel_ps.so")
    blob = open(path, "rb").read()
    assert b"oracle_" not in blob and b"liboracle" not in blob
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kubeflow-controller_b200")):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".go")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in src.lower(), f"{f} mentions the oracle"


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="box has a GPU")
def test_compute_fails_loudly_without_gpu(pkg):
    """No CPU fallback: on a machine without a CUDA device the apply is an error."""
    import numpy as np
    ctx = pkg.Context(1)
    var = np.zeros(16, np.float32)
    g = np.zeros(16, np.float32)
    arr = (ctypes.c_void_p * 1)(g.ctypes.data)
    hp = pkg.HParams()
    rc = pkg.load_library().accel_apply_sgd(ctx.handle, 0, var.ctypes.data, arr, 1, 0, 16,
                                            ctypes.byref(hp), None, None)
    assert rc == -3        # ACCEL_ERR_CUDA
    assert b"CUDA" in pkg.load_library().accel_last_error()
    assert not var.any()


def test_argument_errors_without_touching_a_device(pkg):
    L = pkg.load_library()
    assert L.accel_ctx_create(None, 1, None) == -1
    assert L.accel_adam_advance(None) == -1
    assert L.accel_ctx_gpu_of(None, 0) == -1
    ctx = pkg.Context(1)
    with pytest.raises(pkg.AccelError):
        ctx.set_kernel(kernel=9)
    with pytest.raises(pkg.AccelError):
        ctx.set_kernel(threads=100)
    with pytest.raises(pkg.AccelError):
        ctx.set_kernel(unroll=3)
    ctx.set_kernel(pkg.KERNEL_LDG256, 256, 2, 2)
    buf = ctypes.create_string_buffer(4)
    assert L.accel_tfjob_service_key(b"a-long-job-name", 0, 0, buf, 4) == -5   # BUFSIZE
    assert b"bytes" in L.accel_last_error()


def test_go_wrapper_binds_only_declared_symbols():
    """The cgo wrapper cannot be compiled here (no Go toolchain); at least every C symbol, type
    and constant it names must exist in include/accel_ps.h."""
    header = open(os.path.join(ROOT, "include", "accel_ps.h")).read()
    go_dir = os.path.join(ROOT, "kubeflow-controller_b200", "go", "pkg", "accel")
    used = set()
    for f in os.listdir(go_dir):
        if f.endswith(".go"):
            used |= set(re.findall(r"\bC\.((?:accel|ACCEL)_[A-Za-z0-9_]+)", open(os.path.join(go_dir, f)).read()))
    assert used, "no C references found in the Go wrapper"
    missing = sorted(u for u in used if not re.search(r"\b" + re.escape(u) + r"\b", header))
    assert not missing, f"Go wrapper references undeclared C names: {missing}"
    # and call arity: every C.accel_xxx(...) call passes as many arguments as the prototype has
    protos = {}
    flat = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    for m in re.finditer(r"\b(accel_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", flat, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    for f in os.listdir(go_dir):
        if not f.endswith(".go"):
            continue
        src = open(os.path.join(go_dir, f)).read()
        for m in re.finditer(r"C\.(accel_[a-z0-9_]+)\(", src):
            if m.group(1) not in protos:      # a type conversion such as C.accel_stream_t(...)
                continue
            depth, i, commas = 1, m.end(), 0
            while depth:
                c = src[i]
                depth += c in "([{"
                depth -= c in ")]}"
                commas += (c == "," and depth == 1)
                i += 1
            inner = src[m.end():i - 1].strip()
            n_args = 0 if not inner else commas + 1
            assert n_args == protos[m.group(1)], (
                f"{f}: C.{m.group(1)} called with {n_args} args, header declares {protos[m.group(1)]}")
