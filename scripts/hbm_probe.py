"""What elementwise streaming reaches on this GPU (context for the k_drift roofline): torch copy / add on
arrays the size of the particle arrays, CUDA events, 20 repetitions after warm-up."""
import torch
n = 16777216
dev = "cuda:0"
def t(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for words, label in ((4, "float4 arrays (268 MB)"), (16, "1 GB arrays")):
    a = torch.empty(n * words, dtype=torch.float32, device=dev).normal_()
    b = torch.empty_like(a).normal_(); c = torch.empty_like(a)
    by = a.numel() * 4
    ms = t(lambda: c.copy_(a)); print(label, "copy  %.3f ms  %.0f GB/s" % (ms, 2 * by / ms / 1e6))
    ms = t(lambda: torch.add(a, b, out=c)); print(label, "add   %.3f ms  %.0f GB/s" % (ms, 3 * by / ms / 1e6))
    ms = t(lambda: a.mul_(1.0001)); print(label, "scale in place %.3f ms  %.0f GB/s" % (ms, 2 * by / ms / 1e6))
# many streams at once: 10 in-place scales of separate 268 MB arrays in one foreach call
xs = [torch.empty(n * 4, dtype=torch.float32, device=dev).normal_() for _ in range(10)]
ms = t(lambda: torch._foreach_mul_(xs, 1.0001)); print("foreach in-place scale of 10 x 268 MB %.3f ms  %.0f GB/s" % (ms, 20 * n * 16 / ms / 1e6))
