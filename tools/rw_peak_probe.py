"""Pure-write, pure-read and copy bandwidth of this B200 (torch ops over 8 GiB), to put the write-dominated
forward kernels (12.2 GB written / 0.6 GB read per launch) next to the right ceiling."""
import torch
n = 2 * 1024 ** 3  # f32 elements = 8 GiB
a = torch.empty(n, dtype=torch.float32, device="cuda")
b = torch.empty(n, dtype=torch.float32, device="cuda")
def t(fn, reps=5):
  fn(); torch.cuda.synchronize()
  best = 1e9
  for _ in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
  return best
gb = n * 4 / 1e9
ms = t(lambda: a.fill_(1.0)); print("write (fill_): %.3f ms  %.0f GB/s" % (ms, gb / ms * 1e3))
ms = t(lambda: a.zero_()); print("write (memset): %.3f ms  %.0f GB/s" % (ms, gb / ms * 1e3))
ms = t(lambda: a.sum()); print("read (sum): %.3f ms  %.0f GB/s" % (ms, gb / ms * 1e3))
ms = t(lambda: b.copy_(a)); print("copy: %.3f ms  %.0f GB/s read+write" % (ms, 2 * gb / ms * 1e3))
