"""Index arithmetic of the MaCow CUDA kernels checked on the CPU: csrc/macow_kernel.cuh compiled by g++ as a
one-thread-per-CTA emulation (tests/emulation/macow_emul.cpp) against the oracle and the reference fixtures.
This proves tap offsets, weight transposes, tile order and stage sequencing; thread mapping, barriers and
shuffles are only exercised by the GPU tests (tests/test_gpu_macow.py)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import macow_oracle as mo

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emulation", "macow_emul.cpp")
OUT = os.path.join(HERE, "emulation", "_build", "libmacow_emul.so")
KERNEL = os.path.join(os.path.dirname(HERE), "ipoke-main-opticalflow_b200", "csrc", "macow_kernel.cuh")
ACT = {"relu": 0, "elu": 1, "leaky_relu": 2}
ORDER = ["conv1", "conv2", "conv3", "conv4"]


@pytest.fixture(scope="module")
def emul():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(SRC), os.path.getmtime(KERNEL)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-Wno-attributes",
                        "-o", OUT, SRC], check=True)
    L = ctypes.CDLL(OUT)
    L.emul_unit_floats.restype = ctypes.c_longlong
    return L


def pack(L, sd, C, Hd, k0, k1):
    arrs = []
    for n in ORDER:
        arrs += [sd[n + ".net.shift_conv.weight"], sd[n + ".net.conv1x1.conv.bias"],
                 sd[n + ".net.conv1x1.conv.weight_g"], sd[n + ".net.conv1x1.conv.weight_v"]]
    arrs += [sd["actnorm1.log_scale"], sd["actnorm1.bias"], sd["actnorm2.log_scale"], sd["actnorm2.bias"]]
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in arrs]
    ptrs = (ctypes.c_void_p * 20)(*[a.ctypes.data for a in arrs])
    packed = np.zeros(L.emul_unit_floats(C, Hd, k0, k1), dtype=np.float32)
    L.emul_pack(ptrs, packed.ctypes.data_as(ctypes.c_void_p), C, Hd, k0, k1)
    return packed


def run(L, packed, x, meta, s0, s1, reverse, tile=1, n_units=1, logdet0=None, qc=4, ksplit=1, ws=0):
    C, Hd, k0, k1, act, alpha = meta
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, _, H, W = x.shape
    out = np.empty_like(x)
    ld = np.zeros(B, dtype=np.float32) if logdet0 is None else logdet0.astype(np.float32).copy()
    L.emul_run(packed.ctypes.data_as(ctypes.c_void_p), n_units, x.ctypes.data_as(ctypes.c_void_p),
               out.ctypes.data_as(ctypes.c_void_p), ld.ctypes.data_as(ctypes.c_void_p), int(logdet0 is not None), B,
               C, Hd, H, W, k0, k1, ACT[act], ctypes.c_float(alpha), s0, s1, tile, int(reverse), qc, ksplit, ws)
    return out, ld


def load(name):
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    sd = {k[4:]: g[k] for k in g.files if k.startswith("sd::")}
    C, k0, k1 = (int(v) for v in g["meta"][:3])
    Hd = sd["conv1.net.shift_conv.weight"].shape[0]
    return g, sd, (C, Hd, k0, k1, str(g["activation"]), float(g["alpha"]))


@pytest.mark.parametrize("name", ["macow_elu", "macow_relu", "macow_lrelu"])
def test_emulated_kernel_matches_reference_fixtures(emul, name):
    g, sd, meta = load(name)
    packed = pack(emul, sd, *meta[:4])
    for tile in (1, 2, 3, 64):                                   # forward tiling must not change the result
        out, ld = run(emul, packed, g["x"], meta, 0, 6, False, tile=tile, qc=8 if tile == 2 else 4, ws=tile & 1)
        np.testing.assert_allclose(out, g["out64"], rtol=2e-5, atol=2e-5)
        np.testing.assert_allclose(ld, g["logdet64"], rtol=2e-5, atol=3e-4)
    for qc, ksplit in ((4, 1), (8, 2), (8, 3)):                  # 3 does not divide the channel counts: ragged last group
        x_smp, _ = run(emul, packed, g["z_in"], meta, 0, 6, True, qc=qc, ksplit=ksplit, ws=ksplit & 1)
        np.testing.assert_allclose(x_smp, g["x_smp"], rtol=1e-4, atol=1e-4)
    x_rec, _ = run(emul, packed, g["out"], meta, 0, 6, True)
    np.testing.assert_allclose(x_rec, g["x"], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("name", ["macow_elu", "macow_relu"])
def test_emulated_single_stages(emul, name):
    g, sd, meta = load(name)
    packed = pack(emul, sd, *meta[:4])
    act, alpha = meta[4], meta[5]
    for s, stage in enumerate(mo.STAGES):
        want, want_ld = mo.unit_forward(sd, g["x"], act, alpha, np.float64, stages=(stage,))
        out, ld = run(emul, packed, g["x"], meta, s, s + 1, False, tile=2)
        np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-5, err_msg=stage)
        np.testing.assert_allclose(ld, want_ld, rtol=1e-5, atol=1e-4, err_msg=stage)
        back = mo.unit_reverse(sd, g["z_in"], act, alpha, np.float64, stages=(stage,))
        got, _ = run(emul, packed, g["z_in"], meta, s, s + 1, True)
        np.testing.assert_allclose(got, back, rtol=1e-5, atol=1e-5, err_msg=stage)
    # logdet accumulation and a stage sub-range
    out, ld = run(emul, packed, g["x"], meta, 1, 4, False, tile=1, logdet0=np.full(g["x"].shape[0], 2.5))
    want, want_ld = mo.unit_forward(sd, g["x"], act, alpha, np.float64, stages=mo.STAGES[1:4])
    np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ld, want_ld + 2.5, rtol=1e-5, atol=1e-4)


def random_unit(rng, C, k0, k1, gain=0.3):
    """A synthetic MaCowUnit state dict with the reference's key names and shapes."""
    Hd = 4 * C if C <= 96 else min(2 * C, 512)
    sd = {}
    for n in ORDER:
        ks = (k0, k1) if n in ("conv1", "conv2") else (k1, k0)
        sd[n + ".net.shift_conv.weight"] = (rng.standard_normal((Hd, C) + ks) / np.sqrt(C * k0 * k1)).astype(np.float32)
        sd[n + ".net.conv1x1.conv.bias"] = (0.1 * rng.standard_normal(2 * C)).astype(np.float32)
        sd[n + ".net.conv1x1.conv.weight_g"] = (gain * (0.5 + rng.random((2 * C, 1, 1, 1)))).astype(np.float32)
        sd[n + ".net.conv1x1.conv.weight_v"] = (0.05 * rng.standard_normal((2 * C, Hd, 1, 1))).astype(np.float32)
    for n in ("actnorm1", "actnorm2"):
        sd[n + ".log_scale"] = (0.1 * rng.standard_normal((C, 1, 1))).astype(np.float32)
        sd[n + ".bias"] = (0.1 * rng.standard_normal((C, 1, 1))).astype(np.float32)
    return sd, Hd


@pytest.mark.parametrize("C,k0,k1,H,W,act", [(8, 3, 5, 5, 7, "elu"), (32, 2, 3, 8, 8, "relu"), (3, 1, 1, 4, 9, "leaky_relu"),
                                                (4, 2, 5, 8, 12, "elu")])      # float4 line reads with 5 centred taps
def test_emulated_two_unit_chain_other_kernels(emul, C, k0, k1, H, W, act):
    rng = np.random.default_rng(C)
    (sd0, Hd), (sd1, _) = random_unit(rng, C, k0, k1), random_unit(rng, C, k0, k1)
    meta = (C, Hd, k0, k1, act, 0.8)
    packed = np.concatenate([pack(emul, sd0, C, Hd, k0, k1), pack(emul, sd1, C, Hd, k0, k1)])
    x = rng.standard_normal((2, C, H, W)).astype(np.float32)
    m, ld0 = mo.unit_forward(sd0, x, act, 0.8)
    want, ld1 = mo.unit_forward(sd1, m, act, 0.8)
    out, ld = run(emul, packed, x, meta, 0, 12, False, tile=3, n_units=2, qc=8)
    np.testing.assert_allclose(out, want, rtol=3e-5, atol=3e-5)
    np.testing.assert_allclose(ld, ld0 + ld1, rtol=3e-5, atol=3e-4)
    back, _ = run(emul, packed, out, meta, 0, 12, True, n_units=2)
    np.testing.assert_allclose(back, x, rtol=2e-4, atol=2e-4)
    z = rng.standard_normal((2, C, H, W)).astype(np.float32)
    want_rev = mo.unit_reverse(sd0, mo.unit_reverse(sd1, z, act, 0.8), act, 0.8)
    got, _ = run(emul, packed, z, meta, 0, 12, True, n_units=2, ksplit=min(C, 4), ws=1)       # odd C: unaligned unit offsets
    np.testing.assert_allclose(got, want_rev, rtol=2e-4, atol=2e-4)
