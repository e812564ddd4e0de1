import sys
import torch
import conftest  # noqa: F401
import test_deform_gpu as T

def main():
    for N in (96, 128, 129, 300, 2048, 2049, 4992, 5037):
        for rep_i in range(2):
            rec = T.oracle_record(N=N, W=128)
            rep = {}
            try:
                T.check(rec, *T.run_product(rec), report=rep)
                st = "PASS"
            except AssertionError:
                st = "FAIL"
            worst = sorted(rep.items(), key=lambda kv: -kv[1])[:4]
            print(N, rep_i, st, [(k, f"{v:.2e}") for k, v in worst])
            sys.stdout.flush()

main()
