#!/usr/bin/env python
"""bench.py -- Panda collision checks/sec (FK + self [+ env]) on B200, one process per GPU.

A *step* is one pass of the hot path (batched FK + broadphase + narrowphase + verdict pack) over one
batch of synthetic configurations; `value` = checks/s with q resident in HBM, `e2e` = the same through
the C-ABI with pinned HOST buffers (H2D of q and D2H of verdict/pair bits inside the timed region).
`--impl reference` times the CPU restatement of the reference path (oracle/, all host threads) on a
bounded sample of the same workload.  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "panda_collision_checks_per_sec"
UNIT = "checks/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--batch", type=int, default=0, help="configurations per GPU per step (0 = workload default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--inflight", type=int, default=2, help="blocking C-ABI calls kept in flight by the device-resident leg (value)")
    ap.add_argument("--e2e-inflight", type=int, default=2, help="blocking C-ABI calls kept in flight by the end-to-end leg")
    ap.add_argument("--max-rounds", type=int, default=20000, help="c4: cap on lock-step RRTConnect rounds per step")
    ap.add_argument("--connect-steps", type=int, default=8, help="c4: connect-walk steps validated per lock-step round")
    ap.add_argument("--plan-time", type=float, default=0.5,
                    help="c4: OMPL `time` budget in seconds.  The reference arm gives EVERY query 1.0 s on its own host thread "
                         "(Planner.plan's default); the batched call applies its budget to the whole lock-step batch.  0.5 s is "
                         "the default because it already solves more of the queries than the reference does with 1.0 s per query "
                         "(solve_rate is printed in both lines); 1.0 and 0.25 are recorded under profiles/")
    return ap.parse_args()


def workload_config(args):
    from tests import workloads as WL

    return WL.get(args.workload, args.batch)


class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md 'clocks' line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_leg(desc, q, threads, budget_s, chunk):
    """Time the oracle on successive chunks of q until ~budget_s of wall time is spent."""
    from oracle import oracle as O

    tw = time.perf_counter()
    O.collide_batch(desc, q[:1000], threads=threads)  # warm-up (page-in, OpenMP pool); also sizes the chunks
    tw = max(time.perf_counter() - tw, 1e-6)
    chunk = int(max(1000, min(chunk, 1000 * budget_s / tw / 4)))  # at most about a quarter of the budget per chunk
    done, t0 = 0, time.perf_counter()
    while done < q.shape[0] and time.perf_counter() - t0 < budget_s:
        m = min(chunk, q.shape[0] - done)
        O.collide_batch(desc, q[done:done + m], threads=threads)
        done += m
    dt = time.perf_counter() - t0
    return done / dt, done, dt


PLAN_METRIC = "panda_rrtconnect_plans_per_sec"
PLAN_UNIT = "plans/s"


def run_c4_reference(args):
    """Reference arm of configs[3]: the oracle's RRTConnect restatement (OMPL semantics, one query per host thread,
    `time` = 1 s per query as Planner.plan passes) on a bounded sample of the same seeded queries."""
    from oracle import oracle as O
    from mplib_b200.pymp import OMPLPlanner
    from tests import workloads as WL

    wl = WL.get("c4", args.batch)
    desc, w = wl["desc"], wl["world"]
    threads = os.cpu_count() or 1
    pl = OMPLPlanner(w)
    lo, hi = pl.lower_joint_limits, pl.upper_joint_limits
    n_q = max(threads * 2, 16)
    checker = lambda q: O.collide_batch(desc, q, threads=threads)[0]
    per_step, solved_tot = [], 0
    for s in range(args.warmup + args.steps):
        starts, goals = WL.c4_queries(n_q, 5 + 104729 * s, checker)
        t0 = time.perf_counter()
        st, paths, gc, sc = O.rrtconnect_batch(desc, lo, hi, starts, goals, range_=0.1, time_limit=1.0, seed=s, threads=threads)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            per_step.append((int((st == 1).sum()), dt, n_q))
    n_solved = sum(a for a, _, _ in per_step)
    t_tot = sum(b for _, b, _ in per_step)
    val = n_solved / t_tot
    line = {
        "impl": "reference", "metric": PLAN_METRIC, "value": val, "unit": PLAN_UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / max(len(per_step), 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": wl["config"],
        "solve_rate": n_solved / max(sum(c for _, _, c in per_step), 1),
        "cpu_baseline": {"value": val, "unit": PLAN_UNIT, "cores": threads, "kind": "port",
                         "sample": f"{n_q} queries/step of the same seeded workload, one query per thread, OpenMP x{threads}, "
                                   "time limit 1 s per query"},
        "e2e": {"value": val, "unit": PLAN_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_c4(args):
    """configs[3]: plans/s of lock-step RRTConnect.  `value` = solved queries / device time of mpb_rrtconnect_batch
    (CUDA events inside the call); `e2e` = solved / wall time of OMPLPlanner.plan_batch (start/goal validity checks, goal
    images, H2D of the queries, D2H of status + paths)."""
    import torch
    import torch.distributed as dist

    from mplib_b200 import _lib
    from mplib_b200.pymp import OMPLPlanner
    from tests import workloads as WL

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if _lib.load().mpb_device_count() == 0:
        raise SystemExit("bench.py: no CUDA device visible; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # F plan_batch calls in flight (F host threads, each with its own world / scene handle): a lock-step round is a
    # chain of small kernels, so the rounds of independent query batches interleave on the GPU.
    F = max(1, args.inflight)
    wls = [WL.get("c4", args.batch) for _ in range(F)]
    wl = wls[0]
    desc, Q = wl["desc"], wl["batch"]
    pls = []
    for x in wls:
        x["world"].device = local  # the world compiles its scene image on this rank's GPU
        pls.append(OMPLPlanner(x["world"]))
    n_sets = min(args.warmup + args.steps, 4)
    sets = [WL.c4_queries(Q, 5 + 104729 * k + 1000003 * rank, wl["world"].collide_batch) for k in range(n_sets)]
    kw = dict(range=0.1, time=args.plan_time, max_rounds=args.max_rounds, tree_capacity=4096, max_path=2048,
              connect_steps=args.connect_steps)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(k0, k1):
        out, errors = {}, []

        def worker(t):
            try:
                torch.cuda.set_device(local)
                for k in range(k0 + t, k1, F):
                    out[k] = pls[t].plan_batch(*sets[k % n_sets], seed=k, **kw)[2]
            except BaseException as e:
                errors.append(e)

        th = [threading.Thread(target=worker, args=(t,)) for t in range(F)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        if errors:
            raise errors[0]
        return out

    run(0, max(args.warmup, F))
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    solved = launches = rounds = states = 0
    t0 = time.perf_counter()
    res = run(args.warmup, args.warmup + args.steps)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    # device time of the region: the calls of one host thread run back to back, the threads run concurrently
    gpu_ms = max(sum(st["gpu_ms"] for k, st in res.items() if (k - args.warmup) % F == t) for t in range(F))
    for st in res.values():
        solved += st["n_solved"]; launches += st["kernel_launches"]; rounds += st["rounds"]; states += st["states_checked"]
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([gpu_ms, wall], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gpu_ms, wall = float(t[0]), float(t[1])
        c = torch.tensor([solved, launches, states], device="cuda", dtype=torch.float64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        solved, launches, states = float(c[0]), float(c[1]), float(c[2])
    if rank == 0:
        K = args.steps
        f = wl["flop_model"]
        peak_probe = _lib.fp32_peak_tflops(local)
        F_state = f["F_fk"] + 12.0 * desc.n_pairs  # FK + sphere cull per checked state (narrowphase not counted here)
        achieved = states * F_state / (gpu_ms * 1e-3) / 1e12
        line = {
            "metric": PLAN_METRIC, "value": solved / (gpu_ms * 1e-3), "unit": PLAN_UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": gpu_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": wl["config"],
            "solve_rate": solved / (float(Q) * world * K), "plan_time_budget_s": args.plan_time, "calls_in_flight": F,
            "e2e": {"value": solved / wall, "unit": PLAN_UNIT, "h2d_bytes_per_step": int(Q * 2 * desc.state_dim * 4 + Q * 4),
                    "d2h_bytes_per_step": int(Q * 8 + Q * kw["max_path"] * desc.state_dim * 4)},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "fp32", "scope": "all kernels of the planning rounds; FK + cull flops of the checked states only",
                         "achieved": achieved, "peak": peak_probe, "unit": "TFLOP/s", "frac": achieved / peak_probe if peak_probe else None,
                         "traffic": None, "states_checked_per_step": states / K / world, "rounds_per_step": rounds / K},
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload == "c4":
        return run_c4_reference(args)
    wl = workload_config(args)
    desc = wl["desc"]
    threads = os.cpu_count() or 1
    q = wl["make_q"](0, min(wl["batch"], 1 << 18)).astype(np.float64)
    per_step = []
    for s in range(args.warmup + args.steps):
        v, n, dt = cpu_leg(desc, q, threads, budget_s=3.0, chunk=q.shape[0])
        if s >= args.warmup:
            per_step.append((n, dt))
    n_tot = sum(n for n, _ in per_step)
    t_tot = sum(dt for _, dt in per_step)
    val = n_tot / t_tot
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / max(len(per_step), 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": wl["config"],
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{per_step[0][0] if per_step else 0} states/step of the same seeded workload, OpenMP x{threads}"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def ncu_traffic(workload, kernel):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of `kernel` from the committed
    `ncu --set full` capture of this workload (profiles/r01_final_ncu_full_<w>_raw.csv), or None."""
    import csv
    path = os.path.join(ROOT, "profiles", "r01_final_ncu_full_%s_raw.csv" % workload)
    try:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        ik, ir, iw, ig = (hdr.index(c) for c in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size"))
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for r in rows[2:]:
            if kernel in r[ik]:
                b = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
                states = int(float(r[ig])) * 128
                return {"dram_bytes_per_launch": b, "states_per_launch": states, "dram_bytes_per_state": b / states,
                        "source": os.path.relpath(path, ROOT)}
    except Exception:
        pass
    return None


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "c4":
        return run_c4(args)

    import torch
    import torch.distributed as dist

    from mplib_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if _lib.load().mpb_device_count() == 0:
        raise SystemExit("bench.py: no CUDA device visible; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    wl = workload_config(args)
    desc, N, D = wl["desc"], wl["batch"], wl["desc"].state_dim
    dev = _lib.DeviceScene(desc, local)
    W = dev.pair_words
    stream = torch.cuda.current_stream()
    # rotating pool of input batches larger than L2 (126 MB) so no step re-reads cached q
    n_pool = max(2, int(np.ceil(160e6 / (N * D * 4))))
    n_pool = min(n_pool, 16)
    pool_h = [torch.from_numpy(wl["make_q"](rank * 1000 + k, N).astype(np.float32)).pin_memory() for k in range(n_pool)]
    pool_d = [t.cuda(non_blocking=True) for t in pool_h]
    vwords = (N + 31) // 32
    # F calls in flight (F host threads, each with its own scene handle, stream and result set), the way a caller
    # pipelines a blocking API: while one call synchronises to read its counters and the host enqueues its next pass,
    # the other call's kernels keep the GPU busy.  The verdict gather of step k (NCCL, side stream, issued in step
    # order by the main thread) overlaps the kernels of the following steps.  --inflight 1 is the strictly serial figure.
    F = max(1, args.inflight)
    devs_d = [dev] + [_lib.DeviceScene(desc, local) for _ in range(F - 1)]
    streams = [torch.cuda.Stream() for _ in range(F)]
    # verdict gather (NCCL): one all_gather per G steps of a call slot (G x 131 KB for C2), two buffers per slot, so a
    # worker only waits for the gather issued two of its groups ago
    G = 4 if world > 1 else 1
    verdict_d = [[torch.zeros((G, vwords), dtype=torch.int32, device="cuda") for _ in range(2)] for _ in range(F)]
    pairs_d = [[torch.zeros((N, max(W, 1)), dtype=torch.int32, device="cuda") for _ in range(2)] for _ in range(F)]
    gathered = [[torch.zeros((world, G, vwords), dtype=torch.int32, device="cuda") for _ in range(2)] for _ in range(F)] if world > 1 else None
    comm = torch.cuda.Stream() if world > 1 else None
    ev_done = [[torch.cuda.Event() for _ in range(2)] for _ in range(F)]
    ev_gathered = [[torch.cuda.Event() for _ in range(2)] for _ in range(F)]
    verdict_h = torch.zeros(vwords, dtype=torch.int32).pin_memory()
    pairs_h = torch.zeros((N, max(W, 1)), dtype=torch.int32).pin_memory()
    torch.cuda.synchronize()

    def run_device(k0, k1, start_event=None, F=F):
        """Steps k0..k1-1 with F calls in flight; returns the per-step stats in step order."""
        n_of = [len(range(k0 + t, k1, F)) for t in range(F)]      # steps of worker t
        groups = [(n_of[t] + G - 1) // G for t in range(F)]         # gather groups of worker t
        done = {(t, g): threading.Event() for t in range(F) for g in range(groups[t])}
        issued = {(t, g): threading.Event() for t in range(F) for g in range(groups[t])}
        stats = {}
        errors = []

        def worker(t):
            try:
                torch.cuda.set_device(local)
                if start_event is not None:
                    streams[t].wait_event(start_event)
                for j, k in enumerate(range(k0 + t, k1, F)):
                    g, r = j // G, j % G
                    b = g & 1
                    if world > 1 and r == 0 and g >= 2:
                        issued[(t, g - 2)].wait()
                        streams[t].wait_event(ev_gathered[t][b])  # the gather that read this buffer two groups ago
                    stats[k] = devs_d[t].collide_raw(pool_d[k % n_pool], N, verdict_d[t][b][r], pairs_d[t][b], _lib.MPB_DEVICE,
                                                     streams[t].cuda_stream)
                    if r == G - 1 or j == n_of[t] - 1:
                        ev_done[t][b].record(streams[t])
                        done[(t, g)].set()
            except BaseException as e:  # surface worker failures instead of hanging the main thread
                errors.append(e)
                for ev in list(done.values()) + list(issued.values()):
                    ev.set()

        th = [threading.Thread(target=worker, args=(t,)) for t in range(F)]
        for x in th:
            x.start()
        if world > 1:
            for g in range(max(groups)):
                for t in range(F):
                    if g >= groups[t] or errors:
                        continue
                    done[(t, g)].wait()
                    if errors:
                        break
                    b = g & 1
                    with torch.cuda.stream(comm):
                        comm.wait_event(ev_done[t][b])
                        dist.all_gather_into_tensor(gathered[t][b], verdict_d[t][b])  # the only inter-GPU traffic: verdict words
                        ev_gathered[t][b].record(comm)
                    issued[(t, g)].set()
        for x in th:
            x.join()
        if errors:
            raise errors[0]
        for t in range(F):
            stream.wait_stream(streams[t])
        if world > 1:
            stream.wait_stream(comm)  # the last gathers are inside the timed region
        return [stats[k] for k in range(k0, k1)]

    def step_host(k):
        return dev.collide_raw(pool_h[k % n_pool], N, verdict_h, pairs_h, _lib.MPB_HOST, stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run_device(0, args.warmup)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    agg = {"gpu_ms_cull": 0.0, "gpu_ms_narrow": 0.0, "gpu_ms_fk": 0.0, "pairs_after_cull": 0, "pairs_after_spheres": 0, "narrow_iterations": 0, "narrow_capped": 0,
           "kernel_launches": 0}
    e0.record(stream)
    per_step = run_device(args.warmup, args.warmup + args.steps, start_event=e0)
    e1.record(stream)  # the worker streams and the gather stream have been joined into this one
    barrier()
    ms = e0.elapsed_time(e1)
    # per-kernel durations: with several calls in flight the kernels of different calls share the SMs and each call's
    # event intervals stretch, so the kernel split (and the fk_table_kernel roofline) comes from a short serial leg
    # of the same steps, one call in flight, still inside the clock-sampled region
    if F > 1:
        per_step_serial = run_device(args.warmup + args.steps, args.warmup + args.steps + min(args.steps, 5), F=1)
    else:
        per_step_serial = per_step
    clocks = sampler.stop() if rank == 0 else None
    Ks = len(per_step_serial)
    for st in per_step_serial:
        for key in ("gpu_ms_cull", "gpu_ms_narrow", "gpu_ms_fk"):
            agg[key] += st[key] * args.steps / Ks  # scaled to K steps so the per-step divisions below stay uniform
    for st in per_step:
        for key in ("pairs_after_cull", "pairs_after_spheres", "narrow_iterations", "narrow_capped", "kernel_launches"):
            agg[key] += st[key]
        fk_bytes = st["fk_bytes_per_state"]
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())

    # end-to-end through the C-ABI with pinned host buffers.  Two calls are kept in flight (two host threads, each
    # with its own scene handle, streams and pinned result buffers), the way a caller pipelines a blocking API: the
    # H2D copy of one call runs under the kernels of the other.  --e2e-inflight 1 gives the strictly serial figure.
    n_fl = max(1, args.e2e_inflight)
    devs = (devs_d + [_lib.DeviceScene(desc, local) for _ in range(max(0, n_fl - len(devs_d)))])[:n_fl]
    outs = [(verdict_h, pairs_h)] + [(torch.zeros(vwords, dtype=torch.int32).pin_memory(),
                                      torch.zeros((N, max(W, 1)), dtype=torch.int32).pin_memory()) for _ in range(n_fl - 1)]

    def host_worker(t, k0, k1):
        for k in range(k0 + t, k1, n_fl):
            devs[t].collide_raw(pool_h[k % n_pool], N, outs[t][0], outs[t][1], _lib.MPB_HOST, None)

    def run_host(k0, k1):
        th = [threading.Thread(target=host_worker, args=(t, k0, k1)) for t in range(n_fl)]
        for x in th:
            x.start()
        for x in th:
            x.join()

    run_host(0, max(min(args.warmup, 3), n_fl))
    barrier()
    t0 = time.perf_counter()
    run_host(args.warmup, args.warmup + args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    if rank == 0:
        K = args.steps
        total_checks = float(N) * world * K
        value = total_checks / (ms * 1e-3)
        # algorithmic flop model of SURVEY 8(d)
        f = wl["flop_model"]
        iters_per_step = agg["narrow_iterations"] / K
        F_step = N * (f["F_fk"] + 12.0 * desc.n_pairs) + iters_per_step * (2.0 * (5.0 * f["V_mean"] + 33.0) + 150.0)
        kernels_ms = ms / K  # the whole device-timed step (with calls in flight the per-call kernel intervals overlap)
        peak_probe = _lib.fp32_peak_tflops(local)
        props = torch.cuda.get_device_properties(local)
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        peak_formula = props.multi_processor_count * 128 * 2 * sm_mhz * 1e6 / 1e12
        achieved = F_step / (kernels_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        B_alg = 4.0 * D + 4.0
        hbm_achieved = N * B_alg / (kernels_ms * 1e-3) / 1e9
        dominant = "fk_table+sphere+dense" if agg["gpu_ms_cull"] >= agg["gpu_ms_narrow"] else "narrow(+expand)"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": wl["config"],
            "e2e": {"value": total_checks / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(N * D * 4),
                    "d2h_bytes_per_step": int(vwords * 4 + N * max(W, 1) * 4), "calls_in_flight": n_fl},
            "calls_in_flight": F,
            "gpu_launches": int(agg["kernel_launches"]),
            "clocks": clocks,
            "roofline": {
                "bound": "fp32", "scope": "all kernels of one step (fk_table, sphere, dense, scan, scatter, narrow [, expand, scan, scatter, narrow])",
                "achieved": achieved, "peak": peak_probe, "unit": "TFLOP/s", "frac": achieved / peak_probe if peak_probe else None,
                "peak_source": "in-run FFMA probe (mpb_fp32_peak); formula SMs*128*2*f_sampled = %.1f" % peak_formula,
                "flop_per_check": F_step / N, "traffic": None, "dominant_kernel": dominant,
                "kernel_ms_per_step": {"fk_table+sphere+dense+scan+scatter": agg["gpu_ms_cull"] / K, "narrow(+expand)": agg["gpu_ms_narrow"] / K,
                                       "measured": "serial leg of %d steps, one call in flight" % Ks},
                "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                        "bytes_per_check": B_alg, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                # the one HBM-bound kernel of the pass, timed alone with CUDA events on its own stream inside the C-ABI call
                "fk_table_kernel": {"bound": "hbm", "achieved": N * fk_bytes / (agg["gpu_ms_fk"] / K * 1e-3) / 1e9 if agg["gpu_ms_fk"] > 0 else None,
                                    "peak": hbm_peak, "unit": "GB/s",
                                    "frac": N * fk_bytes / (agg["gpu_ms_fk"] / K * 1e-3) / 1e9 / hbm_peak if agg["gpu_ms_fk"] > 0 else None,
                                    "bytes_per_state": fk_bytes, "ms_per_step": agg["gpu_ms_fk"] / K,
                                    "traffic": ncu_traffic(args.workload, "fk_table_kernel"),
                                    "share_of_step": agg["gpu_ms_fk"] / max(agg["gpu_ms_cull"] + agg["gpu_ms_narrow"], 1e-9)},
            },
            "stats_per_step": {"pairs_after_spheres": agg["pairs_after_spheres"] / K, "pairs_after_cull": agg["pairs_after_cull"] / K, "support_evals": iters_per_step,
                               "capped": agg["narrow_capped"] / K},
        }
        if not args.no_cpu_baseline and world == 1:
            qs = wl["make_q"](0, min(N, 1 << 20)).astype(np.float64)
            v, n, dt = cpu_leg(desc, qs, 1, budget_s=12.0, chunk=50000)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": f"first {n} states of the seeded workload, oracle.c single thread, {dt:.1f}s"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
