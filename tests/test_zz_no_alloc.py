"""arachne_ksw_extend_static (include/arachne_cuda.h): the allocation-free variant of the ksw_extend2 batch, against the existing entry
point arachne_ksw_extend_batch on the same problems, bit for bit, on two calls in a row (the second smaller, with other parameters, so
anything the shared static scratch kept from the first call would show).  Collected last on purpose (file name)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MAX_PROBLEMS = 65536  # ARACHNE_KSW_STATIC_MAX_PROBLEMS


def _static(A, torch, qs, ts, h0, w, end_bonus, zdrop, max_qlen=None, stream=None):
    import util
    q, qo = util.pack(qs); t, to = util.pack(ts)
    n = len(qs)
    dev = "cuda:%d" % A.device
    d = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    # at least one element each so that data_ptr() is never NULL
    dq, dt_ = d(np.concatenate([q, [0]]), np.uint8), d(np.concatenate([t, [0]]), np.uint8)
    dqo, dto, dh, dw = d(qo, np.int64), d(to, np.int64), d(np.concatenate([h0, [1]]), np.int32), d(np.concatenate([w, [1]]), np.int32)
    out = torch.full((n + 1, 6), 12345, dtype=torch.int32, device=dev)
    status = torch.full((1,), 77, dtype=torch.int32, device=dev)
    cells = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    f = A.lib.arachne_ksw_extend_static
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_void_p, C.c_int64] + [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
    mq = max_qlen if max_qlen is not None else max([len(x) for x in qs] + [1])
    rc = f(A.device, C.byref(A.opt), n, dq.data_ptr(), dqo.data_ptr(), dt_.data_ptr(), dto.data_ptr(), dh.data_ptr(), dw.data_ptr(), mq,
           end_bonus, zdrop, out.data_ptr(), status.data_ptr(), cells.data_ptr(), stream)
    if rc:
        A.lib.arachne_last_error.restype = C.c_char_p
        return rc, A.lib.arachne_last_error().decode(), None, None
    torch.cuda.synchronize()
    return 0, out.cpu().numpy()[:n], int(status.item()), int(cells.item())


@pytest.mark.gpu
def test_ksw_extend_static_matches_batch_entry_twice():
    import torch
    import util
    from arachne_b200 import ArachneCuda
    A = ArachneCuda()
    assert A.device_count() > 0
    qs, ts, h0, w = util.extension_problems(211, 6000)
    assert any(int(h0[i]) + len(qs[i]) > 255 for i in range(len(qs))) and any(int(h0[i]) + len(qs[i]) <= 255 for i in range(len(qs)))
    # ---- call 1: the whole set
    exp, exp_cells, _ = A.ksw_extend(*util.pack(qs), *util.pack(ts), h0, w, end_bonus=5, zdrop=100)
    rc, out, status, cells = _static(A, torch, qs, ts, h0, w, 5, 100)
    assert rc == 0 and status == 0, (rc, out)
    assert np.array_equal(out, exp) and cells == exp_cells
    # ---- call 2, straight after: fewer and other problems, other parameters; nothing of call 1 may survive in the shared scratch
    sel = list(range(len(qs) - 1, 0, -3))[:1500]
    qs2 = [qs[i] for i in sel]; ts2 = [ts[i][: max(0, len(ts[i]) - 7)] for i in sel]; h2 = h0[sel]; w2 = np.maximum(1, w[sel] // 2).astype(np.int32)
    exp2, exp_cells2, _ = A.ksw_extend(*util.pack(qs2), *util.pack(ts2), h2, w2, end_bonus=0, zdrop=20)
    rc, out2, status2, cells2 = _static(A, torch, qs2, ts2, h2, w2, 0, 20, max_qlen=255)
    assert rc == 0 and status2 == 0, (rc, out2)
    assert np.array_equal(out2, exp2) and cells2 == exp_cells2
    # ---- and call 1 again on a side stream: same answer as the first time
    s = torch.cuda.Stream()
    rc, out3, status3, cells3 = _static(A, torch, qs, ts, h0, w, 5, 100, stream=s.cuda_stream)
    assert rc == 0 and status3 == 0 and np.array_equal(out3, exp) and cells3 == exp_cells


@pytest.mark.gpu
def test_ksw_extend_static_limits_and_refusals():
    import torch
    import util
    from arachne_b200 import ArachneCuda
    A = ArachneCuda()
    qs, ts, h0, w = util.extension_problems(5, 400)
    # beyond the static scratch: refused up front with a clear message, nothing launched
    big = MAX_PROBLEMS + 1
    z = torch.zeros(8, dtype=torch.int64, device="cuda:%d" % A.device)
    f = A.lib.arachne_ksw_extend_static
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_void_p, C.c_int64] + [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
    p = z.data_ptr()
    assert f(A.device, C.byref(A.opt), big, p, p, p, p, p, p, 150, 5, 100, p, p, None, None) == -1
    A.lib.arachne_last_error.restype = C.c_char_p
    msg = A.lib.arachne_last_error().decode()
    assert str(MAX_PROBLEMS) in msg and "at most" in msg
    assert f(A.device, C.byref(A.opt), 4, p, p, p, p, p, p, 0, 5, 100, p, p, None, None) == -1      # max_qlen out of range
    assert f(A.device, C.byref(A.opt), 4, None, p, p, p, p, p, 150, 5, 100, p, p, None, None) == -1  # null input
    # a problem the kernels cannot take (h0 = 0) is reported through d_status, the others are still answered
    h_bad = h0.copy(); h_bad[17] = 0
    rc, out, status, _ = _static(A, torch, qs, ts, h_bad, w, 5, 100)
    assert rc == 0 and status == 18 and tuple(out[17]) == (-1,) * 6
    keep = [i for i in range(len(qs)) if i != 17]
    exp, _, _ = A.ksw_extend(*util.pack([qs[i] for i in keep]), *util.pack([ts[i] for i in keep]), h0[keep], w[keep], end_bonus=5, zdrop=100)
    assert np.array_equal(out[keep], exp)
    # query longer than the declared bound: refused per problem as well
    lens = [len(x) for x in qs]
    mq = sorted(lens)[len(lens) // 2]
    rc, out, status, _ = _static(A, torch, qs, ts, h0, w, 5, 100, max_qlen=mq)
    first_long = next(i for i, l in enumerate(lens) if l > mq)
    assert rc == 0 and status == first_long + 1
    # n = 0 is a no-op that still clears the status word
    rc, out, status, cells = _static(A, torch, [], [], np.zeros(0, np.int32), np.zeros(0, np.int32), 5, 100, max_qlen=10)
    assert rc == 0 and status == 0 and cells == 0 and out.shape == (0, 6)
