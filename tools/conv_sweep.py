"""Conv microbench sweep (BASELINE.json configs[4]): ResNet-50 / VGG layer shapes, fprop / dgrad / wgrad,
TFLOP/s (2*N*K*C*R*S*P*Q per pass) and fraction of the measured and nominal bf16 peaks.
usage: python tools/conv_sweep.py [out.tsv]"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402

RESNET = [  # (N, H, C, K, R, stride, pad)
    (256, 56, 64, 64, 1, 1, 0), (256, 56, 64, 64, 3, 1, 1), (256, 56, 64, 256, 1, 1, 0), (256, 56, 256, 64, 1, 1, 0),
    (256, 56, 256, 128, 1, 1, 0), (256, 56, 128, 128, 3, 2, 1), (256, 28, 128, 512, 1, 1, 0), (256, 56, 256, 512, 1, 2, 0),
    (256, 28, 512, 128, 1, 1, 0), (256, 28, 128, 128, 3, 1, 1), (256, 28, 512, 256, 1, 1, 0), (256, 28, 256, 256, 3, 2, 1),
    (256, 14, 256, 1024, 1, 1, 0), (256, 28, 512, 1024, 1, 2, 0), (256, 14, 1024, 256, 1, 1, 0), (256, 14, 256, 256, 3, 1, 1),
    (256, 14, 1024, 512, 1, 1, 0), (256, 14, 512, 512, 3, 2, 1), (256, 7, 512, 2048, 1, 1, 0), (256, 14, 1024, 2048, 1, 2, 0),
    (256, 7, 2048, 512, 1, 1, 0), (256, 7, 512, 512, 3, 1, 1),
]
VGG = [(128, 224, 64, 64, 3, 1, 1), (128, 112, 64, 128, 3, 1, 1), (128, 112, 128, 128, 3, 1, 1), (128, 56, 128, 256, 3, 1, 1),
       (128, 56, 256, 256, 3, 1, 1), (128, 28, 256, 512, 3, 1, 1), (128, 28, 512, 512, 3, 1, 1), (128, 14, 512, 512, 3, 1, 1)]


def time_it(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    peak = peaks.get("bf16_tflops", 1590.0)
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
    out.write("net\tshape\tpass\tus\tTFLOP/s\tfrac_measured_peak\tfrac_nominal_2250\n")
    for net, shapes in (("resnet50", RESNET), ("vgg16", VGG)):
        for N, H, C, K, R, stride, pad in shapes:
            x = torch.randn(N, H, H, C, device="cuda").to(torch.bfloat16)
            w = (torch.randn(K, R, R, C, device="cuda") * 0.05).to(torch.bfloat16)
            P = (H + 2 * pad - R) // stride + 1
            dy = torch.randn(N, P, P, K, device="cuda").to(torch.bfloat16)
            fl = 2.0 * N * P * P * K * C * R * R
            g = torch.ones(K, device="cuda"); b0 = torch.zeros(K, device="cuda")
            rm = torch.zeros(K, device="cuda"); rv = torch.ones(K, device="cuda")
            for name, fn in (("fprop", lambda: RF.conv2d_fprop(x, w, stride=stride, pad=pad)),
                             ("fprop+bnstats", lambda: RF.conv2d_fprop_bn_stats(x, w, g, b0, 1e-5, 0.1, rm, rv,
                                                                                stride=stride, pad=pad)),
                             ("dgrad", lambda: RF.conv2d_dgrad(dy, w, tuple(x.shape), stride=stride, pad=pad)),
                             ("wgrad", lambda: RF.conv2d_wgrad(x, dy, tuple(w.shape), stride=stride, pad=pad))):
                ms = time_it(fn)
                tf = fl / ms / 1e9
                out.write(f"{net}\tN{N} {H}x{H} C{C} K{K} {R}x{R} s{stride}\t{name}\t{ms*1e3:.1f}\t{tf:.1f}\t{tf/peak:.3f}\t{tf/2250:.3f}\n")
            del x, w, dy
    if out is not sys.stdout:
        out.close()


if __name__ == "__main__":
    main()
