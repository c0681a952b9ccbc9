#!/usr/bin/env python3
"""Development probe (GPU box): what the host link of this box gives in one direction and in both at once, for contiguous
copies and for the strided copies the all-host commit makes (column strips of a row-major trace with a 5320-byte pitch:
cudaMemcpy2DAsync, host side strided, device side packed).  Answers VERDICT r2 weak #7: is 64 GB/s both ways (29 + 35) the
ceiling of the box or of the library's strip pipeline?  Pinned host memory, CUDA events, best of 3."""
import ctypes
import sys

import torch

rt = None
for name in ("libcudart.so.12", "libcudart.so"):
    try:
        rt = ctypes.CDLL(name)
        break
    except OSError:
        pass
if rt is None:
    import glob, os
    cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    rt = ctypes.CDLL(cands[0])
vp, sz = ctypes.c_void_p, ctypes.c_size_t
rt.cudaMemcpy2DAsync.argtypes = [vp, sz, vp, sz, sz, sz, ctypes.c_int, vp]
rt.cudaMemcpyAsync.argtypes = [vp, vp, sz, ctypes.c_int, vp]
H2D, D2H = 1, 2

PITCH = 5320
rows = (int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20)
host_in = torch.empty(rows * PITCH, dtype=torch.uint8, pin_memory=True)
host_out = torch.empty(rows * PITCH, dtype=torch.uint8, pin_memory=True)
host_in.zero_()
host_out.zero_()
dev_in = torch.empty(rows * PITCH, dtype=torch.uint8, device="cuda")
dev_out = torch.empty(rows * PITCH, dtype=torch.uint8, device="cuda")
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()


def check(rc):
    assert rc == 0, "cuda error %d" % rc


def copy(direction, w, stream):
    """one pass over the matrix: segments of w bytes per row (w == PITCH: one contiguous copy)"""
    st = vp(stream.cuda_stream)
    if w == PITCH:
        if direction == H2D:
            check(rt.cudaMemcpyAsync(dev_in.data_ptr(), host_in.data_ptr(), rows * PITCH, H2D, st))
        else:
            check(rt.cudaMemcpyAsync(host_out.data_ptr(), dev_out.data_ptr(), rows * PITCH, D2H, st))
        return rows * PITCH
    nstrips = PITCH // w
    for k in range(nstrips):
        if direction == H2D:
            check(rt.cudaMemcpy2DAsync(dev_in.data_ptr() + k * rows * w, w, host_in.data_ptr() + k * w, PITCH, w, rows, H2D, st))
        else:
            check(rt.cudaMemcpy2DAsync(host_out.data_ptr() + k * w, PITCH, dev_out.data_ptr() + k * rows * w, w, w, rows, D2H, st))
    return nstrips * rows * w


def timed(fn):
    best = None
    for _ in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s_in.wait_event(e0)
        s_out.wait_event(e0)
        nbytes = fn()
        a, b = torch.cuda.Event(), torch.cuda.Event()
        a.record(s_in)
        b.record(s_out)
        torch.cuda.current_stream().wait_event(a)
        torch.cuda.current_stream().wait_event(b)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None or ms < best else best
    return nbytes, best


print("matrix: %d rows x %d bytes pitch = %.2f GB per direction, pinned host memory" % (rows, PITCH, rows * PITCH / 1e9))
print("%-28s %10s %10s %12s %12s" % ("segment per row", "H2D GB/s", "D2H GB/s", "both: H2D", "both: D2H"))
for w in (PITCH, 128, 384, 896, 1024, 1408, 2656):
    nb, ms_h = timed(lambda: copy(H2D, w, s_in))
    _, ms_d = timed(lambda: copy(D2H, w, s_out))
    _, ms_b = timed(lambda: (copy(H2D, w, s_in), copy(D2H, w, s_out))[0])
    label = "contiguous" if w == PITCH else "%d B (%d columns)" % (w, w // 8)
    print("%-28s %10.1f %10.1f %12.1f %12.1f   (both directions together: %.1f GB/s)" % (label, nb / ms_h / 1e6, nb / ms_d / 1e6, nb / ms_b / 1e6, nb / ms_b / 1e6,
                                                                                     2 * nb / ms_b / 1e6))
