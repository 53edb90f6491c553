"""``lmi_invert_static`` -- the query-to-bucket inversion for callers that cannot allocate device memory.

Some users of the C ABI have no way to hand ``lmi_invert`` its scratch array: a plugin inside a host
application whose memory pool already owns the GPU (a ``cudaMalloc`` from the plugin fails or breaks the
host's budget), a serving thread that must not take the device-wide synchronisation ``cudaMalloc`` /
``cudaFree`` imply, or a foreign-language caller over FFI with no device allocator at hand.  For them the
library carries the scratch itself, as a fixed 4 MiB array in its own image (``LMI_INVERT_STATIC_MAX_INTS``
int32 per GPU), and ``lmi_invert_static`` takes the same arguments as ``lmi_invert`` minus ``workspace``.

What is checked here, through the raw C ABI (ctypes, device pointers, no package code in between):
  * every output array equals ``lmi_invert``'s bit for bit (integer work: the tolerance is zero);
  * on two calls in a row, the second with a different and smaller shape, so whatever the first call left
    in the shared array would show if the kernels relied on it being clean;
  * the call leaves the amount of free device memory untouched;
  * a batch whose scratch need exceeds the built-in array is refused with ``LMI_E_SCRATCH_LIMIT`` and
    nothing is written.

The file sorts last on purpose: ``pytest -x`` stops at the first failure, and a problem in this add-on
must not hide the parity tests.  It passes alone in a fresh process (the first call resolves the array's
device address) and after the rest of the suite.
"""
import numpy as np
import pytest
import torch

LMI_E_SCRATCH_LIMIT = -4
LMI_INVERT_STATIC_MAX_INTS = 1 << 20


def _inputs(rs, nq, B, C, max_rows):
    order = np.stack([rs.permutation(C)[:B] for _ in range(nq)]).astype(np.int32)
    order[rs.rand(nq, B) < 0.02] = -1          # probes the inversion must ignore
    sizes = rs.randint(0, max_rows, C)
    sizes[rs.rand(C) < 0.1] = 0                # buckets without rows on this GPU
    off = np.zeros(C + 1, dtype=np.int64)
    off[1:] = np.cumsum(sizes)
    return order, off, int(sizes.max())


def _outputs(dev, P, C, max_items, fill):
    """The seven caller-owned output arrays, pre-filled so an unwritten cell is visible."""
    mk = lambda n: torch.full((n,), fill, dtype=torch.int32, device=dev)
    return [mk(C), mk(C + 1), mk(C), mk(P), mk(P), mk(max_items * 8), mk(2)]


def _run_both(L, check, rs, nq, B, C, range_len, tail_len, tail_pm):
    dev = torch.device("cuda")
    order, off, biggest = _inputs(rs, nq, B, C, 3000)
    rmax = max(1, -(-biggest // (tail_len or range_len)))
    P = nq * B
    max_items = int(L.lmi_max_items(nq, B, C, rmax))
    need = int(L.lmi_invert_workspace_ints(nq, B, C))
    assert need <= LMI_INVERT_STATIC_MAX_INTS
    d_order = torch.from_numpy(order).to(dev)
    d_off = torch.from_numpy(off).to(dev)

    # the existing entry point, with scratch the caller allocated
    exp = _outputs(dev, P, C, max_items, -7)
    work = torch.empty(need, dtype=torch.int32, device=dev)
    check(L.lmi_invert(d_order.data_ptr(), nq, B, C, d_off.data_ptr(), range_len, tail_len, tail_pm,
                       *[t.data_ptr() for t in exp], max_items, work.data_ptr(), 0), "lmi_invert")
    torch.cuda.synchronize()

    # the new one: same inputs, same outputs, no workspace argument
    got = _outputs(dev, P, C, max_items, -7)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    check(L.lmi_invert_static(d_order.data_ptr(), nq, B, C, d_off.data_ptr(), range_len, tail_len, tail_pm,
                              *[t.data_ptr() for t in got], max_items, 0), "lmi_invert_static")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]

    n_items = int(exp[6][0].item())
    assert n_items > 0
    names = ["counts", "offsets", "bucket_lists", "pair_list", "slot_of_pair", "items", "n_items"]
    for name, e, g in zip(names, exp, got):
        e, g = e.cpu().numpy(), g.cpu().numpy()
        if name == "items":            # cells past n_items are not part of the result
            e, g = e[: n_items * 8], g[: n_items * 8]
        if name == "pair_list":        # filled up to the number of live pairs
            live = int(exp[1][C].item())
            e, g = e[:live], g[:live]
        assert np.array_equal(e, g), f"{name} differs for {(nq, B, C, range_len, tail_len, tail_pm)}"
    return free_before, free_after


@pytest.mark.gpu
def test_invert_static_equals_invert_on_consecutive_calls():
    from learnedmetricindex_b200 import _native

    L = _native.lib()
    rs = np.random.RandomState(11)
    # first call on this device: resolves the array's address (set-up), then a big shape that fills a lot
    # of the shared array ...
    _run_both(L, _native.check, rs, 20000, 16, 50, 4096, 0, 0)
    # ... and straight after it smaller and differently shaped batches that reuse the same cells.  From here
    # on the call only enqueues kernels, so free device memory must not move across it.
    for shape in [(37, 3, 11, 128, 0, 0), (5000, 4, 120, 1024, 256, 300), (10000, 4, 120, 1024, 0, 0),
                  (3000, 7, 1500, 2048, 512, 150)]:
        before, after = _run_both(L, _native.check, rs, *shape)
        assert before == after, f"free device memory moved across lmi_invert_static{shape}: {before} -> {after}"


@pytest.mark.gpu
def test_invert_static_refuses_batches_beyond_its_array():
    from learnedmetricindex_b200 import _native

    L = _native.lib()
    dev = torch.device("cuda")
    nq, B, C = 1 << 20, 1, 257            # 4096 chunks x 257 buckets: just over the built-in array
    assert int(L.lmi_invert_workspace_ints(nq, B, C)) > LMI_INVERT_STATIC_MAX_INTS
    order = torch.zeros(nq * B, dtype=torch.int32, device=dev)
    off = torch.arange(C + 1, dtype=torch.int64, device=dev) * 64
    outs = _outputs(dev, nq * B, C, 16, -7)
    rc = L.lmi_invert_static(order.data_ptr(), nq, B, C, off.data_ptr(), 64, 0, 0,
                             *[t.data_ptr() for t in outs], 16, 0)
    torch.cuda.synchronize()
    assert rc == LMI_E_SCRATCH_LIMIT
    assert b"scratch" in L.lmi_error_string(rc)
    with pytest.raises(_native.LmiError, match="scratch"):
        _native.check(rc, "lmi_invert_static")
    for t in outs:                          # refused before anything was launched
        assert bool((t == -7).all())
    # the same shape one bucket narrower fits exactly
    assert int(L.lmi_invert_workspace_ints(nq, B, 256)) == LMI_INVERT_STATIC_MAX_INTS
