"""GPU: expann_gpu_scan_device_noalloc (include/expann_gpu.h) -- the exact fp32 scan with its scratch memory in the
library's static device array -- against expann_gpu_query_device on the same handle, through the C ABI.

Keys are compared bit for bit (the tolerance of every brute-force test here: none), on calls made in a row on one stream
(the array keeps the last call's contents). Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


def keys_of(call, q_dev, m, k, stream):
    import torch
    out = torch.full((m, k), -1, dtype=torch.int64, device=q_dev.device)
    call(q_dev.data_ptr(), m, k, out.data_ptr(), stream)
    return out


@pytest.mark.parametrize("metric", [0, 1])
def test_noalloc_scan_equals_query_device_twice_in_a_row(ex, metric):
    import torch
    rng = np.random.default_rng(101 + metric)
    base = rng.standard_normal((30_011, 128)).astype(np.float32)
    e = ex.BruteForceEngine(metric=metric)
    e.store_many_vectors(base); e.build()
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    # two different batches in a row, then the first again: every call shares the static array
    batches = [rng.standard_normal((m, 128)).astype(np.float32) for m in (300, 7, 1, 1000)]
    dqs = [torch.from_numpy(b).to(dev) for b in batches]
    for k in (10, 100):
        want = [keys_of(e.query_device, dq, len(b), k, st) for dq, b in zip(dqs, batches)]
        torch.cuda.synchronize()
        outs = [torch.full((len(b), k), -1, dtype=torch.int64, device=dev) for b in batches]
        free0 = None
        for rep in range(3):   # the same calls in a row, no host synchronisation between the calls of a round
            for dq, b, o in zip(dqs, batches, outs):
                e.scan_device_noalloc(dq.data_ptr(), len(b), k, o.data_ptr(), st)
            torch.cuda.synchronize()
            for i, (o, w) in enumerate(zip(outs, want)):
                assert torch.equal(o, w), (metric, k, "round", rep, i)
                o.fill_(-1)
            torch.cuda.synchronize()
            if rep == 0:   # (the first launch of a kernel may load its code: the free-memory check starts behind it)
                free0 = torch.cuda.mem_get_info(dev)[0]
        assert torch.cuda.mem_get_info(dev)[0] == free0, "device memory changed hands during the no-alloc calls"
    e.close()


def test_noalloc_scan_with_id_offset_and_u8_values(ex):
    """a handle with an id offset takes the scratch path even with one row split; u8-valued rows (exact arithmetic)"""
    import torch
    rng = np.random.default_rng(103)
    base = rng.integers(0, 256, (5_000, 64)).astype(np.float32)
    q = rng.integers(0, 256, (2_000, 64)).astype(np.float32)
    e = ex.BruteForceEngine(id_offset=123_456)
    e.store_many_vectors(base); e.build()
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    dq = torch.from_numpy(q).to(dev)
    for m in (2_000, 33):
        want = keys_of(e.query_device, dq, m, 10, st)
        a = keys_of(e.scan_device_noalloc, dq, m, 10, st)
        b = keys_of(e.scan_device_noalloc, dq, m, 10, st)
        torch.cuda.synchronize()
        assert torch.equal(a, want) and torch.equal(b, want), m
        assert int((want[0, 0] & 0xFFFFFFFF).item()) >= 123_456
    e.close()


def test_noalloc_scan_rejects_what_the_static_array_cannot_hold(ex):
    import torch
    from expann_b200._lib import ExpannError, ERR_UNSUPPORTED
    rng = np.random.default_rng(105)
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    e = ex.BruteForceEngine()
    e.store_many_vectors(rng.standard_normal((4_000, 32)).astype(np.float32)); e.build()
    dq = torch.zeros((20_000, 32), dtype=torch.float32, device=dev)
    out = torch.zeros((20_000, 128), dtype=torch.int64, device=dev)
    for m, k, what in ((16_385, 10, "queries"), (10, 129, "k > 128")):
        with pytest.raises(ExpannError) as ei:
            e.scan_device_noalloc(dq.data_ptr(), m, k, out.data_ptr(), st)
        assert ei.value.code == ERR_UNSUPPORTED and what in str(ei.value), (m, k, str(ei.value))
    e.scan_device_noalloc(dq.data_ptr(), 16_384, 10, out.data_ptr(), st)   # the largest batch goes through
    torch.cuda.synchronize()
    e.close()
    e = ex.BruteForceEngine(id_offset=7)            # an id offset sends every list through the scratch array: 16 000 x 128 keys > 2^20
    e.store_many_vectors(rng.standard_normal((4_000, 32)).astype(np.float32)); e.build()
    with pytest.raises(ExpannError) as ei:
        e.scan_device_noalloc(dq.data_ptr(), 16_000, 128, out.data_ptr(), st)
    assert ei.value.code == ERR_UNSUPPORTED and "keys exceed" in str(ei.value), str(ei.value)
    e.scan_device_noalloc(dq.data_ptr(), 8_192, 128, out.data_ptr(), st)    # exactly 2^20 keys
    torch.cuda.synchronize()
    e.close()
    e = ex.BruteForceEngine()                       # d not a multiple of 32: the queries would need a padded copy
    e.store_many_vectors(rng.standard_normal((500, 100)).astype(np.float32)); e.build()
    with pytest.raises(ExpannError) as ei:
        e.scan_device_noalloc(dq.data_ptr(), 4, 10, out.data_ptr(), st)
    assert ei.value.code == ERR_UNSUPPORTED and "multiple of 32" in str(ei.value)
    e.close()
    g = ex.QuantScanEngine(quantizer=0)             # other engines: no such path
    g.store_many_vectors(rng.integers(0, 256, (500, 32)).astype(np.float32)); g.build()
    with pytest.raises(ExpannError) as ei:
        g.scan_device_noalloc(dq.data_ptr(), 4, 10, out.data_ptr(), st)
    assert ei.value.code == ERR_UNSUPPORTED
    g.close()
    mg = ex.BruteForceEngine(devices=[0, 0])
    mg.store_many_vectors(rng.standard_normal((4_000, 32)).astype(np.float32)); mg.build()
    with pytest.raises(ExpannError) as ei:
        mg.scan_device_noalloc(dq.data_ptr(), 4, 10, out.data_ptr(), st)
    assert ei.value.code == ERR_UNSUPPORTED
    mg.close()
