#!/usr/bin/env python3
"""Time the generic NNoM executor on the upstream keyword_spotting example (65,536 device-resident inputs): tensor-core nodes vs the
one-thread-per-output kernels, and check both against each other.  usage: generic_time.py [n_inputs]"""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def main():
    import torch
    import edison_b200 as ed
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    dev = torch.device("cuda", 0)
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    for name in ("nnom_kws_example.bin", "nnom_graph_mnist_simple.bin", "nnom_graph_uci_inception.bin"):
        with ed.Model(os.path.join(ROOT, "tests", "golden", name)) as m:
            x = torch.randint(-128, 128, (n, m.in_size), dtype=torch.int8, device=dev)
            lg = torch.empty(n, m.n_out, dtype=torch.int8, device=dev); sm = torch.empty_like(lg)
            res = {}
            for tag, naive, reps in (("mma", False, 10), ("naive", True, 3)):
                m.set_naive(naive)
                for _ in range(2):
                    chk(L.edison_model_infer(m._h, P(x), n, P(lg), P(sm)))
                t0 = time.perf_counter()
                for _ in range(reps):
                    chk(L.edison_model_infer(m._h, P(x), n, P(lg), P(sm)))
                res[tag] = ((time.perf_counter() - t0) / reps * 1e3, lg.clone())
            print("%-32s %d inputs: tensor-core nodes %.3f ms, one thread per output %.3f ms, x%.2f, same logits %s" % (
                name, n, res["mma"][0], res["naive"][0], res["naive"][0] / res["mma"][0], bool(torch.equal(res["mma"][1], res["naive"][1]))), flush=True)

if __name__ == "__main__":
    main()
