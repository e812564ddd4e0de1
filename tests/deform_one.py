import sys
import torch
import conftest  # noqa: F401
import test_deform_gpu as T
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2049
rec = T.oracle_record(N=N, W=128)
rep = {}
try:
    T.check(rec, *T.run_product(rec), report=rep)
    print("PASS")
except AssertionError:
    print("FAIL")
print(sorted(rep.items(), key=lambda kv: -kv[1])[:5])
