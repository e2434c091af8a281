"""quade_bin_offsets_static (include/quade_b200.h): the bin layout stage with its scratch in a static __device__ array,
for callers that cannot call cudaMalloc.  It has no counterpart in the reference; the check is against the existing
entry point (quade_submit / quade_collect) on the same data, bit-exact, on calls in a row that share the scratch.
Collected last on purpose (a failure here hides nothing after it)."""
import ctypes as C

import numpy as np
import pytest

MAX_PAIRS, MAX_BINS = 1 << 20, 1024


def _layout_from_arenas(r):
    """Per pair, in arena order (= bin order, input order inside a bin): bin, R1 / R2 record bytes and positions,
    read back from what the existing entry point wrote."""
    per_read = []
    for k in range(2):
        a = r.arena(k)
        nl = np.flatnonzero(a == 10)
        assert len(nl) % 4 == 0
        ends = nl[3::4] + 1                                   # one past each record's last newline
        starts = np.concatenate(([0], ends[:-1])).astype(np.uint64) if len(ends) else np.zeros(0, np.uint64)
        bins = np.searchsorted(r.bin_off[k], starts, side="right") - 1
        per_read.append((bins.astype(np.uint32), (ends - starts.astype(np.int64)).astype(np.uint32), starts))
    (b1, l1, p1), (b2, l2, p2) = per_read
    assert (b1 == b2).all()                                   # the i-th record of a bin in R1 pairs with the i-th in R2
    return b1, l1, l2, p1, p2


def _mix(bins, seed):
    """A pair order with the bins interleaved that keeps the input order inside every bin."""
    rng = np.random.default_rng(seed)
    slot = rng.permutation(len(bins))
    dest = np.empty(len(bins), np.int64)
    for b in np.unique(bins):
        idx = np.flatnonzero(bins == b)
        dest[idx] = np.sort(slot[idx])
    order = np.empty(len(bins), np.int64)                     # order[j] = arena index of the pair at mixed position j
    order[dest] = np.arange(len(bins))
    return order


def _existing_entry_point(name, n, first):
    from quade_b200 import DemuxContext
    from quade_b200.synth import Synth
    s = Synth.from_config(name)
    ctx = DemuxContext(s.dual, s.spans(), s.min_qual, s.sample_keys(), max_stream_bytes=64 << 20, max_pairs=n + 16)
    r = ctx.run(s.generate(first, n), last=True)
    assert r.n_pairs == n
    bins, l1, l2, p1, p2 = _layout_from_arenas(r)
    order = _mix(bins, seed=first + 1)
    case = dict(nbins=r.nbins, bin_off=np.stack(r.bin_off), counters=r.counters.copy(), bin=bins[order], len1=l1[order],
                len2=l2[order], off1=p1[order], off2=p2[order])
    ctx.close()
    return case


def _call_static(L, torch, case, want_offsets=True):
    dev = torch.device("cuda:0")
    n, nb = len(case["bin"]), case["nbins"]
    t = lambda a: torch.from_numpy(a.view(np.int32).copy()).to(dev)
    d_bin, d_l1, d_l2 = t(case["bin"]), t(case["len1"]), t(case["len2"])
    # outputs start as garbage: the entry point must write every element it owns
    d_boff = torch.full((2, nb + 1), -7, dtype=torch.int64, device=dev)
    d_cnt = torch.full((nb,), -7, dtype=torch.int64, device=dev)
    d_o1 = torch.full((max(n, 1),), -7, dtype=torch.int64, device=dev)
    d_o2 = torch.full((max(n, 1),), -7, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = L.quade_bin_offsets_static(d_bin.data_ptr(), d_l1.data_ptr(), d_l2.data_ptr(), n, nb, d_boff.data_ptr(), d_cnt.data_ptr(),
                                    d_o1.data_ptr() if want_offsets else None, d_o2.data_ptr() if want_offsets else None,
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.quade_last_error().decode()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return (d_boff.cpu().numpy().astype(np.uint64), d_cnt.cpu().numpy(), d_o1.cpu().numpy()[:n].astype(np.uint64),
            d_o2.cpu().numpy()[:n].astype(np.uint64), free_before - free_after)


def _check(got, case, tag):
    boff, cnt, o1, o2, _ = got
    assert (boff == case["bin_off"]).all(), tag                # bit-exact: integer work
    assert (cnt == case["counters"]).all(), tag
    assert (o1 == case["off1"]).all(), tag
    assert (o2 == case["off2"]).all(), tag


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_existing_one_on_calls_in_a_row():
    import torch
    from quade_b200 import _lib
    L = _lib.lib()
    big = _existing_entry_point("cfg3", 150_000, first=11)       # 769 bins, 37 histogram blocks
    small = _existing_entry_point("cfg2", 20_000, first=3)       # 193 bins, 5 blocks: another row stride in the scratch
    assert big["nbins"] <= MAX_BINS and small["nbins"] != big["nbins"]
    _check(_call_static(L, torch, big), big, "first call")
    second = _call_static(L, torch, small)                       # the scratch still holds the first call's rows
    _check(second, small, "second call")
    third = _call_static(L, torch, big)
    _check(third, big, "third call")
    assert second[4] == 0 and third[4] == 0, "device memory in use grew across a call: something allocated"
    # offsets not wanted; no pairs at all
    boff, cnt, o1, _, _ = _call_static(L, torch, big, want_offsets=False)
    assert (boff == big["bin_off"]).all() and (cnt == big["counters"]).all() and (o1.astype(np.int64) == -7).all()
    empty = dict(big, bin=big["bin"][:0], len1=big["len1"][:0], len2=big["len2"][:0])
    boff, cnt, _, _, _ = _call_static(L, torch, empty)
    assert not boff.any() and not cnt.any()


def test_static_scratch_entry_point_refuses_sizes_beyond_the_array():
    """Argument checks come before any CUDA call: they run without a device."""
    from quade_b200 import _lib
    L = _lib.lib()
    one = C.c_void_p(256)
    assert L.quade_bin_offsets_static(one, one, one, MAX_PAIRS + 1, 8, one, one, None, None, None) == _lib.ERR_CAPACITY
    assert b"at most 1048576 pairs" in L.quade_last_error()
    assert L.quade_bin_offsets_static(one, one, one, 10, MAX_BINS + 1, one, one, None, None, None) == _lib.ERR_CAPACITY
    assert b"1..1024 bins" in L.quade_last_error()
    assert L.quade_bin_offsets_static(one, one, one, 10, 0, one, one, None, None, None) == _lib.ERR_CAPACITY
    assert L.quade_bin_offsets_static(one, one, one, 10, 8, None, one, None, None, None) == _lib.ERR_ARG
    assert L.quade_bin_offsets_static(one, one, one, 10, 8, one, one, one, None, None) == _lib.ERR_ARG
