"""Per-trip latency floor of the power-iteration launch at small sizes (development tool).
Run alone (N=1) or under torch.distributed.run (N>1).  Prints one JSON line per grid size:
    python tools/latency_bench.py [g ...]        # 3D Laplacian g^3
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    comm = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(es.Comm.unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm = es.Comm(rank, world, idt.cpu().numpy().tobytes(), local)
    grids = [int(a) for a in sys.argv[1:]] or [16, 64, 128, 256]
    # LAT_CONFIGS="name:K=V,K=V;name2:K=V" runs every grid under each set of library knobs (read at matrix creation)
    configs = [("default", {})]
    if os.environ.get("LAT_CONFIGS"):
        configs = []
        for item in os.environ["LAT_CONFIGS"].split(";"):
            name, _, kv = item.partition(":")
            configs.append((name, dict(p.split("=") for p in kv.split(",") if p)))
    for cfg_name, cfg_env in configs:
      for k in [k for k in os.environ if k.startswith("EIGSOL_B200_") and k not in ("EIGSOL_B200_EXCHANGE", "EIGSOL_B200_PEER_STATS")]:
          del os.environ[k]
      os.environ.update(cfg_env)
      for g in grids:
          n = g ** 3
          lo, hi = n * rank // world, n * (rank + 1) // world
          rp, ci, v = gen.laplacian_3d(g, lo, hi)
          dm = es.DeviceMatrix.from_csr_rows(n, lo, rp, ci, v, comm)
          x0 = es.random_unit(n, np.float64, seed=1)
          dm.begin(x0, 2 ** 31 - 2, -1.0)
          dm.step(64)
          best = 1e30
          for _ in range(3):
              ms = dm.step(320)
              mine = ms
              if comm is not None:
                  ms = comm.max_float(ms)
              if ms < best:
                  best, best_mine = ms, mine
          info = dm.info()
          if os.environ.get("EIGSOL_B200_PEER_STATS") and world > 1:
              dm.peer_stats()  # reset
              ms1 = dm.step(320)
              ps = dm.peer_stats()
              la = max(ps["control_launches"], 1)
              print(f"# peer_stats cfg={cfg_name} g={g} rank={rank} us/trip={ms1 / 320 * 1e3:.2f} "
                    f"control_wait_us/launch={ps['control_wait_ns'] / la / 1e3:.2f} halo_warps/launch={ps['halo_warps'] / la:.1f} "
                    f"waiting_warps/launch={ps['halo_waiting_warps'] / la:.1f} "
                    f"mean_halo_wait_us={(ps['halo_wait_ns'] / max(ps['halo_waiting_warps'], 1)) / 1e3:.2f} "
                    f"max_halo_wait_us={ps['halo_wait_max_ns'] / 1e3:.2f}", flush=True)
          r = dm.end(want_vector=False)
          if world > 1 and os.environ.get("LAT_PER_RANK"):
              print(f"# cfg={cfg_name} g={g} rank={rank} us={best_mine / 320 * 1e3:.2f}", flush=True)
          if rank == 0:
              print(json.dumps({"g": g, "n": n, "world": world, "us_per_trip": best / 320 * 1e3, "exchange": info["exchange"],
                                "kernel": info["kernel_name"], "cfg": cfg_name,
                                "alg_GBps_per_rank": info["alg_bytes"] / (best / 320 * 1e-3) / 1e9,
                                "lambda": repr(r.eigenvalue)}), flush=True)
          dm.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
