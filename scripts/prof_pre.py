"""Run the input-pipeline kernels once each at batch 128 (for ncu) and time them with CUDA events."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp import ops
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.data.transforms import sample_aug_params
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
img = SyntheticCIFAR10(B, True).images.cuda()
p = sample_aug_params(B, torch.Generator().manual_seed(1)).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timed(fn, n=20):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]
print(f"B={B} preprocess_aug  {timed(lambda: ops.preprocess(img, p, channels=4)):.1f} us")
print(f"B={B} preprocess_plain {timed(lambda: ops.preprocess(img, None, channels=4)):.1f} us")
print(f"B={B} preprocess_col aug   {timed(lambda: ops.preprocess_col(img, p, want_sums=True)):.1f} us")
print(f"B={B} preprocess_col plain {timed(lambda: ops.preprocess_col(img, None, want_sums=True)):.1f} us")
