"""pt_image_preprocess_static — the allocation-free variant of pt_image_preprocess (scratch in a file-scope device
array of 8 MiB instead of a caller workspace; no counterpart in the reference)
— through the C ABI against the existing entry point.  Byte / fixed-point work on the same kernels: the
bar is bit-exact, as in tests/test_image_gpu.py.  Collected last on purpose (file name): a failure here must not hide
another test under `pytest -x`.  GPU only."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MEAN, STD, RESCALE = (0.5, 0.5, 0.5), (0.5, 0.5, 0.5), 1 / 255


def _existing(lib, src, S, dev):
    H, W, _ = src.shape
    lib.pt_image_preprocess_workspace.restype = C.c_longlong
    n = lib.pt_image_preprocess_workspace(H, W, S, S)
    ws = torch.empty(n, dtype=torch.uint8, device=dev)
    out = torch.empty(3, S, S, dtype=torch.float32, device=dev)
    u8 = torch.empty(S, S, 3, dtype=torch.uint8, device=dev)
    rc = lib.pt_image_preprocess(C.c_void_p(src.data_ptr()), H, W, C.c_longlong(src.stride(0)), C.c_void_p(out.data_ptr()),
                                 C.c_void_p(u8.data_ptr()), S, S, *(C.c_float(m) for m in MEAN),
                                 *(C.c_float(s) for s in STD), C.c_double(RESCALE), C.c_void_p(ws.data_ptr()),
                                 C.c_longlong(n), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.pt_last_error()
    return out, u8, n


def _static(lib, src, S, out, u8):
    """The call under test: inputs and outputs only, no workspace, nothing allocated inside."""
    H, W, _ = src.shape
    return lib.pt_image_preprocess_static(C.c_void_p(src.data_ptr()), H, W, C.c_longlong(src.stride(0)),
                                          C.c_void_p(out.data_ptr()), C.c_void_p(u8.data_ptr()), S, S,
                                          *(C.c_float(m) for m in MEAN), *(C.c_float(s) for s in STD),
                                          C.c_double(RESCALE), C.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_static_scratch_entry_point_matches_workspace_entry_point(cuda_dev):
    from projectiontrainer_b200 import _lib
    lib = _lib.lib()
    lib.pt_image_preprocess_static_max_bytes.restype = C.c_longlong
    cap = lib.pt_image_preprocess_static_max_bytes()
    assert cap == 8 << 20
    g = torch.Generator().manual_seed(11)
    S = 384
    # two calls in a row on the shared array, the second with a smaller plan (different tap-table offsets and a
    # shorter intermediate): what the first call left behind must not leak into it; then the first size again
    shapes = [(2500, 2048), (300, 777), (2500, 2048)]
    srcs = [torch.randint(0, 256, (H, W, 3), generator=g, dtype=torch.uint8).to(cuda_dev) for H, W in shapes]
    refs = [_existing(lib, s, S, cuda_dev) for s in srcs]
    assert all(n <= cap for _, _, n in refs)
    outs = [(torch.full((3, S, S), float("nan"), device=cuda_dev), torch.zeros(S, S, 3, dtype=torch.uint8, device=cuda_dev))
            for _ in srcs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(cuda_dev)
    free0, _ = torch.cuda.mem_get_info(cuda_dev)
    for s, (o, u) in zip(srcs, outs):               # back to back on one stream, no synchronisation in between
        assert _static(lib, s, S, o, u) == 0, lib.pt_last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(cuda_dev)
    assert torch.cuda.memory_allocated(cuda_dev) == before and free1 >= free0 - (1 << 20)      # nothing was allocated
    for (o, u), (ro, ru, _) in zip(outs, refs):
        assert torch.equal(u, ru)                   # resized bytes: bit-exact
        assert torch.equal(o, ro)                   # f32 pixel_values: same kernels, same operation order: exact


def test_static_scratch_entry_point_rejects_oversized_plans(cuda_dev):
    from projectiontrainer_b200 import _lib
    lib = _lib.lib()
    lib.pt_image_preprocess_workspace.restype = C.c_longlong
    lib.pt_image_preprocess_static_max_bytes.restype = C.c_longlong
    H, W, S = 9000, 64, 384                         # intermediate 9000 x 384 x 3 bytes > 8 MiB
    assert lib.pt_image_preprocess_workspace(H, W, S, S) > lib.pt_image_preprocess_static_max_bytes()
    src = torch.zeros(H, W, 3, dtype=torch.uint8, device=cuda_dev)
    out = torch.full((3, S, S), 5.0, device=cuda_dev)
    u8 = torch.zeros(S, S, 3, dtype=torch.uint8, device=cuda_dev)
    rc = _static(lib, src, S, out, u8)
    torch.cuda.synchronize()
    assert rc != 0
    msg = lib.pt_last_error().decode()
    assert "pt_image_preprocess_static" in msg and "scratch" in msg and str(8 << 20) in msg, msg
    assert (out == 5.0).all()                       # rejected before any launch
