"""The movement kernels of csrc/move.cu at DeepLab's 16 x 513 x 513 shapes: CUDA-event time and algorithmic GB/s.
python tools/move_profile.py [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch


def main():
    from data_free_quantization_b200 import _lib as L
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    dev = "cuda:0"
    lib = L.lib()
    slot = torch.tensor([-2.0, 3.0, 5.0 / 255.0, 0.0], device=dev)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=dev)
    flush = torch.empty(64 << 20, device=dev)

    def timed(name, fn, nbytes, reps=10):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            ts.append(a.elapsed_time(b))
        ms = sorted(ts)[len(ts) // 2]
        print(f"{name:44s} {ms:7.4f} ms  {nbytes / ms / 1e6:8.1f} GB/s (algorithmic, L2 flushed)")

    for C, H, OH, st in ((256, 33, 129, True), (21, 129, 513, False), (256, 1, 33, True), (256, 32, 128, True),
                         (21, 128, 512, False)):
        x = torch.randn(N, C, H, H, device=dev)
        y = torch.empty(N, C, OH, OH, device=dev)
        args = (x.data_ptr(), y.data_ptr(), N, C, H, H, OH, OH, 1, slot.data_ptr(), 8, stats.data_ptr() if st else None)
        timed(f"qinterp {C}x{H}x{H} -> {OH}x{OH}", lambda: L.check(lib.dfq_qinterp_bilinear_f32(*args, L.stream()), "i"),
              4 * (x.numel() + y.numel()))
    for C, H in ((256, 129), (48, 129), (256, 33)):
        x = torch.randn(N, C, H, H, device=dev)
        y = torch.empty(N, C + 48, H, H, device=dev)
        cnt = C * H * H
        args = (x.data_ptr(), y.data_ptr(), N, cnt, cnt, (C + 48) * H * H, slot.data_ptr(), 8, stats.data_ptr())
        timed(f"qcopy (cat operand) {C}x{H}x{H}", lambda: L.check(lib.dfq_qcopy_f32(*args, L.stream()), "c"), 8 * x.numel())
    for C, H, p in ((32, 257, 1), (144, 129, 1), (384, 33, 2), (960, 33, 4)):
        x = torch.randn(N, C, H, H, device=dev)
        y = torch.empty(N, C, H + 2 * p, H + 2 * p, device=dev)
        args = (x.data_ptr(), y.data_ptr(), N, C, H, H, p, p, p, p, stats.data_ptr())
        timed(f"pad2d {C}x{H}x{H} +{p}", lambda: L.check(lib.dfq_pad2d_f32(*args, L.stream()), "p"),
              4 * (x.numel() + y.numel()))


if __name__ == "__main__":
    main()
