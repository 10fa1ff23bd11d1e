"""GPU: lime_choose_static -- clusterChoose with its scratch in a static device array, no allocation inside the
call (no counterpart in the reference) -- against lime_choose on the
same matrices, byte for byte, on two calls in a row that share the scratch array; sizes beyond the maximum are
refused.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import corpus as C_

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

MAX_READS = 16777216          # LIME_CHOOSE_STATIC_MAX_READS (include/lime_cuda.h)


def _static_choose(lib, torch, sim, norm, beta, slack=8):
    """-> (bin bytes, pos bytes, n_records host, n_records device) of lime_choose_static on a copy of `sim`."""
    from lime_b200.api import PAIR_SIM_DT
    n_reads, n_refs = sim.shape
    pitch = (n_refs + 15) // 16 * 16
    padded = np.zeros((n_reads, pitch), dtype=np.uint8)
    padded[:, :n_refs] = sim
    d_sim = torch.from_numpy(padded).to("cuda:0")
    cap = 1 + n_reads + int(np.count_nonzero(sim)) + slack
    d_bin = torch.full((cap * 2,), -1, dtype=torch.int32, device="cuda:0")       # {f32 sim, u32 idRef} records
    d_pos = torch.full((max(n_reads, 1),), -1, dtype=torch.int64, device="cuda:0")
    d_n = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    n = C.c_uint64(0)
    err = C.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.lime_choose_static(C.c_void_p(d_sim.data_ptr()), pitch, n_reads, n_refs, norm, C.c_float(beta),
                                C.c_void_p(d_bin.data_ptr()), cap, C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_n.data_ptr()),
                                C.byref(n), C.c_void_p(stream), err, 256)
    assert rc == 0, err.value
    torch.cuda.synchronize()
    assert np.array_equal(d_sim.cpu().numpy(), padded)                          # the matrix is an input
    k = int(n.value)
    recs = d_bin.cpu().numpy().view(PAIR_SIM_DT)
    assert (d_bin[2 * k:] == -1).all()                                           # nothing past the records
    return recs[:k].tobytes(), d_pos.cpu().numpy()[:n_reads].tobytes(), k, int(d_n.item())


def test_choose_static_equals_lime_choose_twice_in_a_row(oracle):
    import torch
    from lime_b200 import _lib
    from lime_b200.api import LimeContext
    lib = _lib.load()
    ctx = LimeContext(0)
    try:
        # two different shapes back to back: a narrow-row kernel on 9,000 reads, then a wide-row kernel on fewer
        # reads (the second call sees the first one's look-back words beyond its own tiles and must not care)
        for n, n_reads, n_refs, alpha, ebwt_mode, beta in ((900_000, 9_000, 40, 16, 1, 0.25), (400_000, 3_001, 300, 12, 0, 0.0)):
            case = C_.synth_case(f"static{n_refs}", n, n_reads, n_refs, seed=401 + n_refs, p_long=2e-5, max_seg=30000)
            ctx.load(case["lcp"], case["da"], case["ebwt"] if ebwt_mode else None, n_reads, n_refs)
            ctx.cluster_similarity(alpha, ebwt_mode, False)
            sim = ctx.sim_matrix().copy()
            norm = (case["read_len"] + 1 - alpha) & 0xFFFFFFFF
            n_rec = ctx.choose(0, norm, beta)
            bin_g, pos_g = ctx.choose_result(0)
            for rep in range(2):
                b, p, k, k_dev = _static_choose(lib, torch, sim, norm, beta)
                assert k == k_dev == n_rec
                assert b == bin_g.tobytes() and p == pos_g.tobytes()
            bin_o, pos_o = oracle.choose(sim, norm, np.float32(beta))
            assert b == bin_o.tobytes() and p == pos_o.tobytes()
    finally:
        ctx.close()


def test_choose_static_refuses_what_it_cannot_hold():
    import torch
    from lime_b200 import _lib
    lib = _lib.load()
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(64, dtype=torch.int64, device="cuda:0")
    n = C.c_uint64(0)
    err = C.create_string_buffer(256)
    ptr = C.c_void_p(buf.data_ptr())
    o = C.c_void_p(out.data_ptr())
    # beyond the static scratch
    rc = lib.lime_choose_static(ptr, 16, MAX_READS + 1, 4, 85, C.c_float(0.25), o, 8, o, None, C.byref(n), None, err, 256)
    assert rc == -1 and b"maximum" in err.value and str(MAX_READS).encode() in err.value
    # bad pitch / no place for the record count
    assert lib.lime_choose_static(ptr, 10, 16, 4, 85, C.c_float(0.25), o, 8, o, None, C.byref(n), None, err, 256) == -1
    assert lib.lime_choose_static(ptr, 16, 16, 4, 85, C.c_float(0.25), o, 8, o, None, None, None, err, 256) == -1
    # a record buffer that is too small is reported, nothing is written past it
    sim = np.zeros((16, 16), dtype=np.uint8)
    sim[:, :4] = 200
    d_sim = torch.from_numpy(sim).to("cuda:0")
    out.fill_(-1)
    pos = torch.zeros(16, dtype=torch.int64, device="cuda:0")
    rc = lib.lime_choose_static(C.c_void_p(d_sim.data_ptr()), 16, 16, 4, 85, C.c_float(0.25), o, 8, C.c_void_p(pos.data_ptr()),
                                None, C.byref(n), None, err, 256)
    torch.cuda.synchronize()
    assert rc == -6 and int(n.value) == 1 + 16 * 5 and (out[8:] == -1).all()
    # no reads: the sentinel alone
    out.fill_(-1)
    rc = lib.lime_choose_static(None, 16, 0, 4, 85, C.c_float(0.25), o, 8, None, None, C.byref(n), None, err, 256)
    torch.cuda.synchronize()
    assert rc == 0 and int(n.value) == 1 and int(out[0]) == 0 and int(out[1]) == -1
