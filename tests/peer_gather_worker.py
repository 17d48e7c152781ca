"""One rank of tests/test_multi_gpu.py::test_peer_gather_*: evaluates its slice of a batch through omo_peer_gather_eval and writes
the gathered f of every call to <dir>/out<rank>.npy. Handles travel through files in <dir> (the test needs no torch).
    python peer_gather_worker.py <dir> <rank> <world> <device> <B> [absent]
With `absent`, the last rank connects and then never evaluates: the other ranks' waits must time out with an error, not hang."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    d, rank, world, device, B = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    absent = len(sys.argv) > 6 and sys.argv[6] == "absent"
    from conftest import load_scene
    from optimo_b200 import synthetic
    from optimo_b200.core import Core, PeerGather, shard_range

    def exchange(handle):
        with open(os.path.join(d, f"h{rank}.tmp"), "wb") as fh:
            fh.write(handle)
        os.rename(os.path.join(d, f"h{rank}.tmp"), os.path.join(d, f"h{rank}.bin"))
        out = []
        t0 = time.time()
        for r in range(world):
            p = os.path.join(d, f"h{r}.bin")
            while not os.path.exists(p):
                if time.time() - t0 > 120:
                    raise RuntimeError(f"rank {rank}: no handle from rank {r}")
                time.sleep(0.01)
            with open(p, "rb") as fh:
                out.append(fh.read())
        return out

    c = Core(synthetic.stick_240(load_scene("stick-figure")), device=device)
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    eng.set_batch_hint(B)                      # the kernel the WHOLE batch gets, whatever this rank's share
    pg = PeerGather(eng, rank, world, B, exchange)
    if absent:
        from optimo_b200.core import EngineError
        if rank == world - 1:                  # mapped, never delivers; stays alive until the others have given up
            t0 = time.time()
            while not all(os.path.exists(os.path.join(d, f"gave_up{r}")) for r in range(world - 1)) and time.time() - t0 < 60:
                time.sleep(0.05)
            pg.close()
            return
        X = synthetic.perturbed_candidates(x0, B)
        b0, nb = shard_range(B, rank, world)
        dX = eng.device_alloc(8 * max(nb, 1) * x0.size)
        eng.h2d(dX, X[b0:b0 + nb])
        t0 = time.time()
        try:
            pg.eval(dX, nb, b0, sync=True)
            outcome = "returned without an error"
        except EngineError as e:
            outcome = f"error after {time.time() - t0:.2f} s: {e}"
        with open(os.path.join(d, f"gave_up{rank}"), "w") as fh:
            fh.write(outcome)
        # the context is still usable: an ordinary evaluation of the same rows
        f = eng.eval(X[b0:b0 + nb])
        np.save(os.path.join(d, f"after{rank}.npy"), f)
        eng.device_free(dX)
        pg.close()
        return
    results = []
    for call in range(3):                      # three calls: both halves of the gathered array, and the first one again
        X = synthetic.perturbed_candidates(x0, B, first=100 * call)
        b0, nb = shard_range(B, rank, world)
        if call == 2 and rank == world - 1:
            nb = 0                             # a rank without candidates still takes part in the exchange
        dX = eng.device_alloc(8 * max(nb, 1) * x0.size)
        if nb:
            eng.h2d(dX, X[b0:b0 + nb])
        pg.eval(dX, nb, b0, sync=True)
        results.append(eng.d2h(pg.values(), (B,)))
        eng.device_free(dX)
    np.save(os.path.join(d, f"out{rank}.npy"), np.stack(results))
    pg.close()


if __name__ == "__main__":
    main()
