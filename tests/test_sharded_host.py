"""Host logic of the sharded optimizer step (no GPU): the packed-buffer parameter layout the C ABI reports, and the
chunk / shard arithmetic of ShardedAdam."""
import ctypes

import pytest

from ipoke_inn_b200 import _capi
from ipoke_inn_b200.optim_sharded import plan_pieces


def _layout(C, E, H, depth, n_flows, conditional, precision):
    L = _capi.lib()
    desc = _capi.InnDesc(C=C, E=E, H=H, depth=depth, n_flows=n_flows, conditional=conditional,
                         precision=_capi.PRECISIONS[precision], alpha=0.9, mlp_slope=0.01)
    plan = ctypes.c_void_p()
    _capi.check(L.ipoke_inn_plan_create(ctypes.byref(desc), 0, ctypes.byref(plan)))
    n_lin = L.ipoke_inn_num_linears(plan)
    n = 8 + n_lin + 2 * n_flows
    out = (ctypes.c_longlong * n)()
    _capi.check(L.ipoke_inn_param_layout(plan, out, n))
    assert L.ipoke_inn_param_layout(plan, out, n - 1) == -1            # too small an output array
    packed = L.ipoke_inn_packed_bytes(plan)
    L.ipoke_inn_plan_destroy(plan)
    return list(out), n_lin, packed


def _linear_numels(C, E, H, depth, n_flows, conditional):
    """nn.Linear weight sizes in the order of ipoke_inn_pack: (block, half-step, net s|t, layer)."""
    out = []
    for k in range(n_flows):
        din = (C // 2) if not conditional else (E if k % 2 == 0 else C // 2 + E)
        for _i in range(2):
            for _net in range(2):
                out += [din * H] + [H * H] * depth + [H * (C // 2)]
    return out


@pytest.mark.parametrize("name,C,E,H,depth,n,cond,macs", [
    ("cfgA", 128, 128, 1536, 3, 8, 1, 237_502_464), ("cfgB", 512, 0, 1024, 2, 8, 0, 83_886_080),
    ("mini", 128, 64, 128, 1, 2, 1, None)])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_weight_spans_tile_the_hi_arena(name, C, E, H, depth, n, cond, macs, precision):
    out, n_lin, packed = _layout(C, E, H, depth, n, cond, precision)
    numels = _linear_numels(C, E, H, depth, n, cond)
    assert out[0] == 1 and out[4] == n_lin == len(numels) and out[5] == n
    N = out[2]
    if macs is not None:
        assert N == macs                                               # P_w of SURVEY.md 8d, no padding
    spans = sorted((out[8 + i], out[8 + i] + numels[i]) for i in range(n_lin))
    assert spans[0][0] == 0 and spans[-1][1] == N
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))         # contiguous, no overlap, no gap
    assert out[3] == (2 * N if precision == "fp32" else -1)
    assert packed >= (4 if precision == "fp32" else 2) * N
    blocks = [(out[8 + n_lin + 2 * k], out[8 + n_lin + 2 * k + 1]) for k in range(n)]
    assert blocks[0][0] == 0 and blocks[-1][1] == N and all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
    per_block = len(numels) // n
    for k, (lo, hi) in enumerate(blocks):                              # a block's Linears sit inside the block's range
        for i in range(k * per_block, (k + 1) * per_block):
            assert lo <= out[8 + i] and out[8 + i] + numels[i] <= hi


def test_padded_shapes_are_not_direct():
    out, _, _ = _layout(8, 16, 64, 2, 8, 1, "fp32")                    # cfg S: E and C/2 are padded to 64
    assert out[0] == 0


def test_plan_pieces_cover_each_chunk_exactly_once():
    ranges = [(0, 4096), (4096, 12288), (12288, 16384), (16384, 32768), (32768, 36864)]
    for world in (1, 2, 8):
        owned = []
        for rank in range(world):
            pieces = plan_pieces(ranges, 2, world, rank)
            assert [(c["kb"], c["ke"]) for c in pieces] == [(3, 5), (1, 3), (0, 1)]   # last block first
            owned.append(pieces)
        for ci in range(3):
            c0 = owned[0][ci]
            assert c0["sharded"] == (world > 1)
            cover = sorted((p[ci]["own_lo"], p[ci]["own_lo"] + p[ci]["own_n"]) for p in owned)
            if world > 1:
                assert cover[0][0] == c0["lo"] and cover[-1][1] == c0["hi"]
                assert all(a[1] == b[0] for a, b in zip(cover, cover[1:]))
                assert all(p[ci]["own_n"] % 4 == 0 for p in owned)
            else:
                assert cover == [(c0["lo"], c0["hi"])]


def test_plan_pieces_indivisible_chunk_is_replicated():
    pieces = plan_pieces([(0, 100), (100, 4196)], 1, 8, 3)
    assert pieces[0]["sharded"] and pieces[0]["own_n"] == 512 and pieces[0]["own_lo"] == 100 + 3 * 512
    assert not pieces[1]["sharded"] and (pieces[1]["own_lo"], pieces[1]["own_n"]) == (0, 100)
