"""Per-CTA phase timing of the tcgen05 GEMM / conv kernel (its built-in clock64 probes, ab_gemm_set_debug): where a
latency-bound launch spends its time. Prints medians over CTAs in microseconds at the SM clock sampled by nvidia-smi.

  python scripts/gemm_phases.py
"""
import ctypes as C
import subprocess
import sys

import torch

sys.path.insert(0, ".")
from arbius_b200 import ops  # noqa: E402
from arbius_b200.ops import native  # noqa: E402


def sm_mhz():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True).stdout.strip()
        return float(out.splitlines()[0])
    except Exception:  # noqa: BLE001
        return 1900.0


def run(label, fn, grid_hint=4096):
    lib = native.load()
    lib.ab_gemm_set_debug.argtypes = [C.c_void_p]
    buf = torch.zeros(8 * grid_hint, dtype=torch.int64, device="cuda")
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    lib.ab_gemm_set_debug(buf.data_ptr())
    fn()
    torch.cuda.synchronize()
    lib.ab_gemm_set_debug(None)
    d = buf.view(-1, 8)
    d = d[d[:, 5] > 0].double()
    mhz = sm_mhz()
    med = lambda t: float(t.median()) / mhz
    print(f"{label:46s} {us:7.1f} us/call  ctas={d.shape[0]:4d}  setup {med(d[:,0]):5.2f}  first-operands {med(d[:,1]-d[:,0]):5.2f}  "
          f"mainloop-issue {med(d[:,2]-d[:,1]):6.2f}  acc-ready {med(d[:,3]-d[:,2]):5.2f}  epilogue {med(d[:,4]-d[:,3]):5.2f} "
          f"(first chunk: tmem-ld {med(d[:,6]-d[:,3]):4.2f} math+store {med(d[:,7]-d[:,6]):4.2f})  "
          f"total {med(d[:,5]):6.2f} us @ {mhz:.0f} MHz", flush=True)


def main():
    torch.manual_seed(0)
    for (b, h, w, c, n) in [(2, 6, 48, 768, 768), (2, 48, 48, 768, 768), (2, 12, 96, 384, 384), (2, 3, 24, 1152, 1152)]:
        x = torch.randn(b, h, w, c, device="cuda").half()
        wt = ops.pack_conv_weight((torch.randn(n, c, 3, 3, device="cuda") * 0.02).half())
        bias = torch.randn(n, device="cuda").half()
        for bn in (32, 64, 128, 256):
            run(f"conv {b}x{h}x{w} {c}->{n} bn{bn}", lambda: ops.conv3x3(x, wt, bias, block_n=bn, splitk=1))
    for (m, n, k) in [(576, 768, 6912), (576, 2304, 768)]:
        x = torch.randn(m, k, device="cuda").half()
        wt = (torch.randn(n, k, device="cuda") * 0.02).half()
        for bn in (32, 64, 128):
            run(f"linear {m}x{n}x{k} bn{bn}", lambda: ops.gemm(x, wt, None, block_n=bn, splitk=1))


if __name__ == "__main__":
    main()
