"""Column-sharded wide path on N GPUs (torchrun): every rank holds a column block of X and Y, the ranks
exchange the n x 2r projection sums and the (p+q) x r moment rows once per EM iteration (NCCL inside the
library).  Checked on rank 0 against the CPU oracle run on the gathered data.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_colshard.py [N p q r steps]"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from ppls_b200 import Context, Data, Fit, synth   # noqa: E402


def block(n, rank, world):
    return slice(n * rank // world, n * (rank + 1) // world)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    N, p, q, r, steps = (int(a) for a in sys.argv[1:6]) if len(sys.argv) >= 6 else (500, 6001, 403, 3, 5)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(Context.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    ctx = Context(local, rank, world, bytes(idt.cpu().numpy().tobytes()))
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 3)
    cx, cy = block(p, rank, world), block(q, rank, world)
    d = Data.synthetic_cols(ctx, N, truth, 777, cx, cy)
    init = dict(st, W=st["W"][cx], C=st["C"][cy])
    f = Fit(d, init, r, steps, -np.inf, "SVD")
    f.run()
    res = f.result()
    Xb, Yb = d.export()
    # gather blocks and estimates on rank 0 through the host (gloo-free: all_gather of padded tensors)
    def gather(a, width):
        t = torch.zeros((a.shape[0], width), dtype=torch.float64, device=dev)
        t[:, :a.shape[1]] = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [o.cpu().numpy() for o in out]
    wx, wy = -(-p // world) + 1, -(-q // world) + 1
    Xs, Ys = gather(Xb, wx), gather(Yb, wy)
    Ws, Cs = gather(res["W"].T, wx), gather(res["C"].T, wy)
    ok = True
    if rank == 0:
        from oracle import cbind, ppls_fused as F
        from helpers import relmax
        X = np.concatenate([Xs[g][:, :block(p, g, world).stop - block(p, g, world).start] for g in range(world)], axis=1)
        Y = np.concatenate([Ys[g][:, :block(q, g, world).stop - block(q, g, world).start] for g in range(world)], axis=1)
        W = np.concatenate([Ws[g][:, :block(p, g, world).stop - block(p, g, world).start] for g in range(world)], axis=1).T
        Cm = np.concatenate([Cs[g][:, :block(q, g, world).stop - block(q, g, world).start] for g in range(world)], axis=1).T
        Lo = cbind.load()
        Xo, Yo = cbind.gen_data(Lo, 777, 0, N, truth)
        assert np.array_equal(X, Xo) and np.array_equal(Y, Yo), "column-block generator != row generator"
        o = F.fused_loop(Lo, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                         EMsteps=steps, atol=-np.inf)
        errs = dict(loglik=relmax(res["loglik"], o["loglik"]), W=relmax(W, o["W"]), C=relmax(Cm, o["C"]),
                    B=relmax(res["B"], o["B"]), sigT=relmax(res["sigT"], o["sigT"]), mu_T=relmax(res["mu_T"], o["mu_T"]),
                    Ctt=relmax(res["Ctt"], o["Ctt"]), sigE=abs(res["sigE"] / o["sigE"] - 1))
        ok = errs["loglik"] < 1e-10 and all(v < 1e-8 for v in errs.values())
        print("COLSHARD", "OK" if ok else "FAIL", world, "ranks", (N, p, q, r), {k: float(f"{v:.2e}") for k, v in errs.items()}, flush=True)
    f.close(); d.close()
    dist.barrier()
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.stdout.flush()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
