"""Exchange recovery on 2 ranks (torchrun): a fit that one rank abandons mid-run times out on the other (PPLS_E_NCCL after
PPLS_XCHG_TIMEOUT_MS), and the NEXT fit on the same contexts runs as if nothing had happened (capi.cu::xchg_recover_dist:
the ranks compare epochs / failed flags when a fit is created and reset the exchange state together).
    PPLS_XCHG_TIMEOUT_MS=400 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_recover.py"""
import os, sys, time
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("PPLS_XCHG_TIMEOUT_MS", "400")
from ppls_b200 import Context, Data, Fit, PPLSError, synth   # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for fuse in ("1", "0"):                              # one-launch form and three-kernel form of the exchange
        os.environ["PPLS_FUSE_TAIL"] = fuse
        N, p, q, r, steps = 40001, 301, 77, 3, 6
        truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
        st = synth.start_params(truth, 3)
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(Context.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        ctx = Context(local, rank, world, bytes(idt.cpu().numpy().tobytes()))
        lo, hi = N * rank // world, N * (rank + 1) // world
        d = Data.synthetic(ctx, hi - lo, p, q, truth, seed=99, row0=lo)
        f = Fit(d, st, r, steps, -np.inf, "SVD"); f.run(); res0 = f.result(want_mu=False); f.close()
        # the failure: rank 1 walks away after 3 passes, rank 0 wants all 7
        f = Fit(d, st, r, steps, -np.inf, "SVD")
        failed, t0 = False, time.time()
        try:
            f.iterate(7 if rank == 0 else 3)
            if rank == 0:
                f.result(want_mu=False)
        except PPLSError as e:
            failed = True
            msg = str(e)
        dt = time.time() - t0
        f.close()
        if rank == 0:
            ok = ok and failed and dt < 10.0
            print("RECOVER fuse=" + fuse, "rank 0 saw the failure:", failed, f"after {dt:.2f} s", flush=True)
        dist.barrier()
        # the next fits must be unaffected, twice (the epoch carries on after the reset)
        for rep in range(2):
            f = Fit(d, st, r, steps, -np.inf, "SVD"); f.run(); res = f.result(want_mu=False); f.close()
            same = all(np.array_equal(res[k], res0[k]) for k in ("W", "C", "B", "sigT", "loglik"))
            ok = ok and same
            if rank == 0:
                print("RECOVER fuse=" + fuse, "fit", rep, "after the failure identical to the one before:", same, flush=True)
        d.close()
        dist.barrier()
        ctx.close()
        dist.barrier()
    t = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    ok = bool(t.item() > 0.5)
    if rank == 0:
        print("RECOVER", "OK" if ok else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
