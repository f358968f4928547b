"""Run the tensor-core parity cases one by one and print error statistics (no asserts).

Usage: python tools/diag_tc.py [substring ...]   -- each case runs in this process; a CUDA fault stops it.
"""
import json
import sys
import traceback
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import tc_cases as T  # noqa: E402


def main():
    filt = sys.argv[1:]
    cases = [("gemm", n, lambda n=n: T.run_gemm(n)) for n in T.GEMM_CASES]
    cases += [("fprop", n, lambda n=n: T.run_conv_fprop(n)) for n in T.CONV_CASES]
    cases += [("fprop_epi", n, lambda n=n: T.run_conv_fprop(n, epilogue=True)) for n in ("conv_3x3_p1",)]
    cases += [("dgrad", n, lambda n=n: T.run_conv_dgrad(n)) for n in T.CONV_CASES]
    cases += [("wgrad", n, lambda n=n: T.run_conv_wgrad(n)) for n in T.CONV_CASES]
    cases += [("first", n, lambda n=n: T.run_first_conv(n)) for n in T.FIRST_CASES]
    for kind, name, fn in cases:
        tag = f"{kind}:{name}"
        if filt and not any(f in tag for f in filt):
            continue
        try:
            st = fn()
            verdict = "PASS" if (st["n_bad"] == 0 and st["finite"]) else "FAIL"
            print(verdict, tag, json.dumps(st), flush=True)
        except Exception as e:  # noqa: BLE001
            print("ERROR", tag, repr(e), flush=True)
            traceback.print_exc()
            if "CUDA" in repr(e) or "cuda" in repr(e):
                print("stopping: CUDA context is likely poisoned", flush=True)
                return


if __name__ == "__main__":
    main()
