#!/usr/bin/env python
"""bench.py -- Panda collision checks/sec (FK + self + env) and RRTConnect plans/sec on B200, one process per GPU.

One run measures every workload BASELINE.json's metric names and prints ONE JSON line:

  headline   configs[2] (C3): Panda + grasped box vs 128 primitives + mesh + 100k-point octree, 2^22 states per GPU per step
  "also"     configs[1] (C2, FK + self-collision), configs[4] (C5, two Pandas), configs[3] (C4, RRTConnect plans/s)

A *step* is one pass of the hot path (batched FK + broadphase + narrowphase + verdict / hit-list output) over one batch
of synthetic configurations.  `value` = checks/s with q resident in HBM (device-pointer C-ABI calls enqueued back to
back on one stream, CUDA events around the K steps); `e2e` = the same through the C-ABI with pinned HOST buffers
(H2D of q, D2H of verdict words + sparse hit list inside the timed region).  Every record carries `roofline`,
`cpu_baseline` (the hierarchical CPU restatement of the reference path, oracle/, timed beside it on a bounded sample)
and `parity` (a sample of the same inputs checked against the exhaustive oracle outside the timed region).
`--impl reference` times the CPU restatement with all host threads on the same workloads.  See DESIGN.md "Measurement".
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
PLAN_METRIC = "panda_rrtconnect_plans_per_sec"
PLAN_UNIT = "plans/s"
BAND = 1e-4  # north_star: verdicts may differ only where the reference signed distance is within 1e-4 m of contact


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5"], help="the headline workload")
    ap.add_argument("--also", default="c2,c5,c4", help="comma list of further workloads reported under \"also\" (\"none\" = just the headline)")
    ap.add_argument("--batch", type=int, default=0, help="configurations (c4: queries) per GPU per step of the headline (0 = workload default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--plan-inflight", type=int, default=2, help="c4: plan_batch calls kept in flight (host threads)")
    ap.add_argument("--max-rounds", type=int, default=20000, help="c4: cap on lock-step RRTConnect rounds per step")
    ap.add_argument("--connect-steps", type=int, default=8, help="c4: connect-walk steps validated per lock-step round")
    ap.add_argument("--plan-slots", type=int, default=4096, help="c4: concurrent searches per plan_batch call; a finished query's slot takes the next pending query")
    ap.add_argument("--plan-rounds-per-query", type=int, default=4096,
                    help="c4: the PER-QUERY budget in lock-step rounds (the role of OMPL's `time`; the reference arm gives every "
                         "query 1.0 s of one host thread, Planner.plan's default).  Both arms print solve_rate and "
                         "grow_calls_per_query so the budgets can be compared")
    ap.add_argument("--plan-time", type=float, default=60.0, help="c4: wall-clock safety limit of one plan_batch call (not the budget)")
    return ap.parse_args()


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
        busy = [x for x in sm if x >= 0.5 * max(mx)] if mx else sm  # samples taken while a kernel ran
        return {"sm_mhz": float(np.median(busy if busy else sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------- CPU legs (oracle/)
def cpu_leg(desc, q, threads, budget_s, accel):
    """Time the hierarchical CPU restatement on successive chunks of q until ~budget_s of wall time is spent."""
    from oracle import oracle as O

    tw = time.perf_counter()
    O.collide_batch(desc, q[:256 * threads], threads=threads, accel=accel)  # warm-up (page-in, OpenMP pool); also sizes the chunks
    tw = max(time.perf_counter() - tw, 1e-6)
    chunk = int(max(256 * threads, min(q.shape[0], 256 * threads * budget_s / tw / 4)))  # about a quarter of the budget per chunk
    done, t0 = 0, time.perf_counter()
    while done < q.shape[0] and time.perf_counter() - t0 < budget_s:
        m = min(chunk, q.shape[0] - done)
        O.collide_batch(desc, q[done:done + m], threads=threads, accel=accel)
        done += m
    dt = time.perf_counter() - t0
    return done / dt, done, dt


def cpu_baseline_checks(wl, threads, budget_s):
    from oracle import oracle as O

    desc = wl["desc"]
    accel = O.Accel(desc)
    qs = wl["make_q"](0, 1 << 18).astype(np.float64)
    v, n, dt = cpu_leg(desc, qs, threads, budget_s, accel)
    return {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "algorithm": "bvh",
            "sample": f"first {n} states of the seeded workload in {dt:.1f} s; oracle.c hierarchical leg (AABB-tree descent of "
                      f"meshes / octrees with early exit, edge-walking hull support), OpenMP x{threads}"}


def cpu_plans(wl, threads, n_q, seed_base, steps, time_limit=1.0):
    """The oracle's RRTConnect restatement (OMPL semantics, `time` = 1 s per query as Planner.plan passes), one query per
    host thread at a time, hierarchical validity checker."""
    from mplib_b200 import workloads as WL
    from mplib_b200.pymp import OMPLPlanner
    from oracle import oracle as O

    desc, w = wl["desc"], wl["world"]
    accel = O.Accel(desc)
    pl = OMPLPlanner(w)
    lo, hi = pl.lower_joint_limits, pl.upper_joint_limits
    checker = lambda q: O.collide_batch(desc, q, threads=threads, accel=accel)[0]
    per_step = []
    for s in range(steps):
        starts, goals = WL.c4_queries(n_q, seed_base + 104729 * s, checker)
        t0 = time.perf_counter()
        st, paths, gc, sc = O.rrtconnect_batch(desc, lo, hi, starts, goals, range_=0.1, time_limit=time_limit, seed=s, threads=threads,
                                               accel=accel)
        per_step.append((int((st == 1).sum()), time.perf_counter() - t0, n_q, gc))
    return per_step


# ------------------------------------------------------------------------------------------------- distributed context
class Ctx:
    def __init__(self):
        import torch

        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        import torch

        if self.world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    def reduce(self, vals, op="max"):
        import torch

        if self.world == 1:
            return [float(v) for v in vals]
        import torch.distributed as dist

        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    def close(self):
        if self.world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def ncu_capture(workload):
    """DRAM bytes and executed FP32 flops PER STATE, per kernel and in total, from the committed `ncu --set full` capture of
    this workload (profiles/r02_ncu_full_<w>_raw.csv: the kernels of whole passes; scripts/gpu_evidence.sh makes it,
    scripts/collect_evidence.sh exports it), or {} when there is none.  States in the capture = fk_table_kernel's threads."""
    import csv

    path = os.path.join(ROOT, "profiles", f"r02_ncu_full_{workload}_raw.csv")
    if not os.path.exists(path):
        return {}
    try:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        col = {c: i for i, c in enumerate(hdr)}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        ops = ["smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
               "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum"]
        have_ops = all(o in col for o in ops)
        out, states = {}, 0
        for r in rows[2:]:
            name = r[col["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").strip()
            b = sum(float(r[col[c]]) * scale.get(units[col[c]], 1.0) for c in ("dram__bytes_read.sum", "dram__bytes_write.sum") if c in col)
            e = out.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "fp32_flop": 0.0 if have_ops else None})
            e["launches"] += 1
            e["dram_bytes"] += b
            if have_ops:
                e["fp32_flop"] += float(r[col[ops[0]]]) + float(r[col[ops[1]]]) + 2.0 * float(r[col[ops[2]]])
            if name == "fk_table_kernel":
                states += int(float(r[col["launch__grid_size"]])) * 128
        first = rows[2][col["Kernel Name"]] if len(rows) > 2 else ""
        if not states or out["fk_table_kernel"]["launches"] != 1 or "fk_table_kernel" not in first:
            return {}  # the capture must be the kernels of exactly one pass, FK first
        return {"source": os.path.relpath(path, ROOT), "states_in_capture": states, "kernels": out,
                "dram_bytes_per_state": sum(v["dram_bytes"] for v in out.values()) / states,
                "fp32_flop_per_state": (sum(v["fp32_flop"] for v in out.values()) / states) if have_ops else None,
                "fk_table_dram_bytes_per_state": out["fk_table_kernel"]["dram_bytes"] / states}
    except Exception:
        return {}


# ------------------------------------------------------------------------------------------------- collision workloads
def parity_sample(wl, dev, n_sample):
    """GPU verdicts / pair flags of the first n_sample states of the seeded workload against the exhaustive oracle."""
    from oracle import oracle as O

    desc = dev.desc
    q = wl["make_q"](0, n_sample)
    v, ph, st = dev.collide(q, want_pairs=True, stats=True)
    vh, hits = dev.collide_hits(q)
    dense_from_hits = np.zeros_like(ph)
    dense_from_hits[hits[:, 0], hits[:, 1]] = True
    threads = os.cpu_count() or 1
    ov, oph = O.collide_batch(desc, q.astype(np.float32).astype(np.float64), threads=threads)
    mism = np.argwhere(ph != oph)
    in_band = 0
    for i, p in mism[:256]:
        sd = O.pair_signed_distance(desc, q[i].astype(np.float32).astype(np.float64), int(p))
        in_band += int(np.isfinite(sd) and abs(sd) < BAND)
    return {"states_checked": int(n_sample), "pairs_checked": int(ph.size), "pair_mismatches": int(len(mism)),
            "mismatches_in_1e-4_band": int(in_band), "mismatches_out_of_band": int(min(len(mism), 256) - in_band),
            "verdict_mismatches": int((v != ov).sum()), "hit_list_equals_dense_mask": bool((dense_from_hits == ph).all() and (vh == v).all()),
            "narrow_capped": int(st["narrow_capped"]), "oracle": "oracle.c exhaustive leaf tests, float64, same float32-rounded inputs"}


def latency_record(wl, dev, world_obj):
    """Host-to-host latency of small mpb_collide_batch(MPB_HOST) calls and the cost of scene edits (item: legacy single-state
    callers and CollisionObject.set_transformation / update_point_cloud users)."""
    from mplib_b200 import _lib

    out = {}
    lib = _lib.load()
    import ctypes as C

    for n in (1, 32, 1024):
        q = np.ascontiguousarray(wl["make_q"](3, n).astype(np.float32))
        vb = np.zeros((n + 31) // 32, dtype=np.uint32)
        ts = []
        for it in range(220):
            t0 = time.perf_counter()
            rc = lib.mpb_collide_batch(dev._h, q.ctypes.data, n, vb.ctypes.data, None, None, _lib.MPB_HOST, None)
            ts.append(time.perf_counter() - t0)
            assert rc == 0
        out[f"collide_host_n{n}_us"] = float(np.median(ts[20:]) * 1e6)
    if world_obj is not None:
        try:
            w = world_obj
            c = w._sync()
            name = next(n for n in w.normal_objects if n not in w.attached_bodies and w.normal_objects[n].geometry.gtype in (1, 2, 3, 4))
            obj = w.normal_objects[name]
            ts = []
            for it in range(12):
                pose = np.concatenate([obj.T[:3, 3] + np.array([0.0, 0.0, 0.001]), [1.0, 0.0, 0.0, 0.0]])
                t0 = time.perf_counter()
                obj.set_transformation(pose)
                w._sync()
                ts.append(time.perf_counter() - t0)
            out["obstacle_move_resync_ms"] = float(np.median(ts[2:]) * 1e3)
            if "cloud" in w.normal_objects:
                from mplib_b200.workloads import box_surface_points

                rng = np.random.Generator(np.random.PCG64(99))
                pts = box_surface_points(rng, 100000, (0.8, 0.8, 0.05), (0.56, 0, -0.025))
                t0 = time.perf_counter()
                w.remove_normal_object("cloud")
                w.add_point_cloud("cloud", pts, 0.01)
                w._sync()
                out["replace_100k_point_cloud_resync_ms"] = float((time.perf_counter() - t0) * 1e3)
        except Exception as e:  # a scene without such an object: report, do not fail the bench
            out["scene_edit_error"] = repr(e)
    return out


def run_checks(name, args, ctx, batch, steps, warmup, headline):
    """One collision workload -> record (rank 0) or None."""
    import torch
    import torch.distributed as dist

    from mplib_b200 import _lib
    from mplib_b200 import workloads as WL

    world, rank, local = ctx.world, ctx.rank, ctx.local
    t_setup = time.perf_counter()
    wl = WL.get(name, batch)
    desc, N, D = wl["desc"], wl["batch"], wl["desc"].state_dim
    dev = _lib.DeviceScene(desc, local)
    t_setup = time.perf_counter() - t_setup
    K, Wm = steps, warmup
    vwords = (N + 31) // 32
    hit_cap = int(N * wl.get("hit_cap_per_state", 2))
    # rotating pool of input batches larger than L2 (126 MB) so no step re-reads cached q
    n_pool = min(max(2, int(np.ceil(160e6 / (N * D * 4)))), 16)
    pool_h = [torch.from_numpy(wl["make_q"](rank * 1000 + k, N).astype(np.float32)).pin_memory() for k in range(n_pool)]
    pool_d = [t.cuda(non_blocking=True) for t in pool_h]
    verdict_d = torch.zeros((K, vwords), dtype=torch.int32, device="cuda")
    hits_d = torch.zeros((hit_cap, 2), dtype=torch.int32, device="cuda")
    nhits_d = torch.zeros(K + Wm + 8, dtype=torch.int32, device="cuda")
    gathered = torch.zeros((world, K, vwords), dtype=torch.int32, device="cuda") if world > 1 else None
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()

    def enqueue(k, slot):
        return dev.collide_hits_raw(pool_d[k % n_pool], N, verdict_d[slot % K], hits_d, hit_cap, nhits_d[slot:slot + 1], _lib.MPB_DEVICE,
                                    stream.cuda_stream)

    with torch.cuda.stream(stream):
        st0 = enqueue(0, K + Wm)  # first device-pointer call on the scene: blocking, calibrates the pass size
        for k in range(Wm):
            enqueue(k, K + k)
        dev.poll(wait=True)
        ctx.barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_host0 = time.perf_counter()
        e0.record(stream)
        for k in range(K):
            enqueue(Wm + k, k)
        t_enq = time.perf_counter() - t_host0
        if world > 1:  # the only inter-GPU traffic of the path: ONE gather of the packed verdict words of the region
            dist.all_gather_into_tensor(gathered, verdict_d)
        e1.record(stream)
    ctx.barrier()
    ms = e0.elapsed_time(e1)
    agg, done, failed = dev.poll(wait=True)
    assert failed == 0
    nh = nhits_d[:K].cpu().numpy().astype(np.int64)
    # per-kernel split: a short leg of the same steps with the per-pass CUDA events switched on (still in the sampled region)
    dev.set_async_timing(True)
    Ks = min(K, 3)
    with torch.cuda.stream(stream):
        for k in range(Ks):
            enqueue(Wm + K + k, K + k)
    split, _, _ = dev.poll(wait=True)
    dev.set_async_timing(False)
    clocks = sampler.stop() if rank == 0 else None
    (ms,) = ctx.reduce([ms], "max")
    tot = ctx.reduce([agg["narrow_iterations"], agg["kernel_launches"], agg["pairs_after_cull"], agg["pairs_after_spheres"],
                      agg["narrow_capped"], float(nh.sum())], "sum")

    # end to end through the C-ABI with pinned host buffers: H2D of q, kernels, D2H of verdict words + sparse hit list
    verdict_h = torch.zeros(vwords, dtype=torch.int32).pin_memory()
    hits_h = torch.zeros((hit_cap, 2), dtype=torch.int32).pin_memory()
    nh_h = np.zeros(1, dtype=np.uint32)
    e2e_hits = []

    def host_step(k):
        dev.collide_hits_raw(pool_h[k % n_pool], N, verdict_h, hits_h, hit_cap, nh_h, _lib.MPB_HOST, None)
        e2e_hits.append(int(nh_h[0]))

    for k in range(min(Wm, 2)):
        host_step(k)
    e2e_hits.clear()
    ctx.barrier()
    t0 = time.perf_counter()
    for k in range(K):
        host_step(Wm + k)
    e2e_s = time.perf_counter() - t0
    (e2e_s,) = ctx.reduce([e2e_s], "max")

    # for information: the same steps asking for the verdict bit only -- what ValidityCheckerTpl::isValid consumes
    # (src/ompl_planner.h:59-62); the kernels then skip pairs of configurations already known to collide
    Kv = min(K, 5)
    with torch.cuda.stream(stream):
        dev.collide_raw(pool_d[0], N, verdict_d[0], None, _lib.MPB_DEVICE, stream.cuda_stream)
        dev.poll(wait=True)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for k in range(Kv):
            dev.collide_raw(pool_d[(Wm + k) % n_pool], N, verdict_d[k % K], None, _lib.MPB_DEVICE, stream.cuda_stream)
        ev1.record(stream)
    dev.poll(wait=True)
    torch.cuda.synchronize()
    (ms_v,) = ctx.reduce([ev0.elapsed_time(ev1)], "max")

    rec = None
    if rank == 0:
        total_checks = float(N) * world * K
        value = total_checks / (ms * 1e-3)
        f = wl["flop_model"]
        iters_per_step = tot[0] / K / world
        F_step = N * (f["F_fk"] + 12.0 * desc.n_pairs) + iters_per_step * (2.0 * (5.0 * f["V_mean"] + 33.0) + 150.0)
        step_ms = ms / K
        peak_probe = _lib.fp32_peak_tflops(local)
        props = torch.cuda.get_device_properties(local)
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        peak_formula = props.multi_processor_count * 128 * 2 * sm_mhz * 1e6 / 1e12
        achieved = F_step / (step_ms * 1e-3) / 1e12
        peaks = load_peaks()
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        B_alg = 4.0 * D + 4.0
        fk_ms = split["gpu_ms_fk"] / Ks
        cull_ms, narrow_ms = split["gpu_ms_cull"] / Ks, split["gpu_ms_narrow"] / Ks
        fk_bytes = split["fk_bytes_per_state"]
        ncu = ncu_capture(name)
        dram_step = ncu["dram_bytes_per_state"] * N if ncu else None
        flop_exec = ncu["fp32_flop_per_state"] * N if ncu and ncu.get("fp32_flop_per_state") is not None else None
        fk_tr = ncu["fk_table_dram_bytes_per_state"] * N if ncu else None
        d2h = int(vwords * 4 + 4 + 8 * (np.mean(e2e_hits) if e2e_hits else 0))
        rec = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": wl["config"],
            "e2e": {"value": total_checks / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(N * D * 4), "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_s / K, "calls_in_flight": 1,
                    "pcie_gbs_per_gpu": {"h2d": N * D * 4 / (e2e_s / K) / 1e9, "d2h": d2h / (e2e_s / K) / 1e9}},
            "calls_in_flight": 1, "host_enqueue_ms_per_step": 1e3 * t_enq / K,
            "verdict_only": {"value": float(N) * world * Kv / (ms_v * 1e-3), "unit": UNIT, "ms_per_step": ms_v / Kv, "steps": Kv,
                             "note": "same steps, verdict bit only (no pair mask / hit list): what isValid needs; not the headline"},
            "gpu_launches": int(tot[1]),
            "clocks": clocks,
            "roofline": {
                "bound": "fp32", "scope": "all kernels of one step (fk_table, sphere, dense, scan, scatter, narrow [, expand, scan, scatter, narrow])",
                "achieved": achieved, "peak": peak_probe, "unit": "TFLOP/s", "frac": achieved / peak_probe if peak_probe else None,
                "peak_source": "in-run FFMA probe (mpb_fp32_peak); formula SMs*128*2*f_sampled = %.1f" % peak_formula,
                "flop_per_check": F_step / N, "traffic": dram_step,
                "executed_fp32_flop_per_step": flop_exec, "executed_over_algorithmic": (flop_exec / F_step) if flop_exec else None,
                "ncu_source": ncu.get("source"),
                "dominant_kernel": "fk_table+sphere+dense" if cull_ms >= narrow_ms else "narrow(+expand)",
                "kernel_ms_per_step": {"fk_table+sphere+dense+scan+scatter": cull_ms, "narrow(+expand)": narrow_ms,
                                       "measured": "leg of %d steps with per-pass CUDA events on the launching stream" % Ks},
                "hbm": {"achieved": N * B_alg / (step_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": N * B_alg / (step_ms * 1e-3) / 1e9 / hbm_peak, "bytes_per_check": B_alg,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                # the one HBM-bound kernel of the pass, timed alone with CUDA events on the launching stream inside the C-ABI call
                "fk_table_kernel": {"bound": "hbm", "achieved": N * fk_bytes / (fk_ms * 1e-3) / 1e9 if fk_ms > 0 else None,
                                    "peak": hbm_peak, "unit": "GB/s",
                                    "frac": N * fk_bytes / (fk_ms * 1e-3) / 1e9 / hbm_peak if fk_ms > 0 else None,
                                    "bytes_per_state": fk_bytes, "ms_per_step": fk_ms,
                                    "traffic": fk_tr,
                                    "share_of_step": fk_ms / max(cull_ms + narrow_ms, 1e-9)},
            },
            "stats_per_step": {"pairs_after_spheres": tot[3] / K / world, "pairs_after_cull": tot[2] / K / world, "support_evals": iters_per_step,
                               "capped": tot[4] / K / world, "hits": tot[5] / K / world},
            "setup_s": {"world_compile_and_scene_create": t_setup, "first_call_blocking_ms": st0.get("host_ms_total")},
        }
        if not args.no_parity:
            rec["parity"] = parity_sample(wl, dev, 4096 if name == "c3" else 8192)
        if headline and not args.no_latency and world == 1:
            rec["latency"] = latency_record(wl, dev, None)
        if not args.no_cpu_baseline and world == 1:
            rec["cpu_baseline"] = cpu_baseline_checks(wl, 1, 6.0)
    del pool_d, verdict_d, hits_d, gathered
    dev.close()
    if rank == 0 and headline and not args.no_latency and world == 1:
        # scene-edit costs on a fresh PlanningWorld of the same workload (the product's host layer drives the C-ABI)
        try:
            w = wl["world"]
            w.device = local
            rec["latency"].update(latency_record(wl, w._sync()["dev"], w))
            w._cache["dev"].close()
        except Exception as e:
            rec["latency"]["scene_edit_error"] = repr(e)
    torch.cuda.empty_cache()
    return rec


# ------------------------------------------------------------------------------------------------- planning workload
def run_c4(args, ctx, batch, steps, warmup):
    """configs[3]: plans/s of lock-step RRTConnect.  `value` = solved queries / device time of mpb_rrtconnect_batch
    (CUDA events inside the call); `e2e` = solved / wall time of OMPLPlanner.plan_batch (start/goal validity checks, goal
    images, H2D of the queries, D2H of status + paths)."""
    import torch

    from mplib_b200 import _lib
    from mplib_b200 import workloads as WL
    from mplib_b200.pymp import OMPLPlanner

    world, rank, local = ctx.world, ctx.rank, ctx.local
    # F plan_batch calls in flight (F host threads, each with its own world / scene handle): a lock-step round is a
    # chain of small kernels, so the rounds of independent query batches interleave on the GPU.
    F = max(1, args.plan_inflight)
    wls = [WL.get("c4", batch) for _ in range(F)]
    wl = wls[0]
    desc, Q = wl["desc"], wl["batch"]
    pls = []
    for x in wls:
        x["world"].device = local  # the world compiles its scene image on this rank's GPU
        pls.append(OMPLPlanner(x["world"]))
    n_sets = min(warmup + steps, 4)
    sets = [WL.c4_queries(Q, 5 + 104729 * k + 1000003 * rank, wl["world"].collide_batch) for k in range(n_sets)]
    kw = dict(range=0.1, time=args.plan_time, max_rounds=args.max_rounds, tree_capacity=4096, max_path=2048,
              connect_steps=args.connect_steps, n_slots=args.plan_slots, max_rounds_per_query=args.plan_rounds_per_query)
    keep = {}

    def run(k0, k1):
        out, errors = {}, []

        def worker(t):
            try:
                torch.cuda.set_device(local)
                for k in range(k0 + t, k1, F):
                    status, paths, st = pls[t].plan_batch(*sets[k % n_sets], seed=k, **kw)
                    out[k] = st
                    keep[k] = (status, paths)
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

    run(0, max(warmup, F))
    ctx.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    solved = launches = rounds = states = grow = 0
    t0 = time.perf_counter()
    res = run(warmup, warmup + steps)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    # device time of the region: the calls of one host thread run back to back, the threads run concurrently
    gpu_ms = max(sum(st["gpu_ms"] for k, st in res.items() if (k - warmup) % F == t) for t in range(F))
    for st in res.values():
        solved += st["n_solved"]; launches += st["kernel_launches"]; rounds += st["rounds"]; states += st["states_checked"]
        grow += st["grow_calls"]
    ctx.barrier()
    clocks = sampler.stop() if rank == 0 else None
    gpu_ms, wall = ctx.reduce([gpu_ms, wall], "max")
    solved, launches, states, grow = ctx.reduce([solved, launches, states, grow], "sum")
    rec = None
    if rank == 0:
        K = steps
        f = wl["flop_model"]
        peak_probe = _lib.fp32_peak_tflops(local)
        F_state = f["F_fk"] + 12.0 * desc.n_pairs  # FK + sphere cull per checked state (narrowphase not counted here)
        achieved = states * F_state / (gpu_ms * 1e-3) / 1e12
        rec = {
            "metric": PLAN_METRIC, "value": solved / (gpu_ms * 1e-3), "unit": PLAN_UNIT, "n_gpus": world, "steps": K,
            "warmup": warmup, "ms_per_step": gpu_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": wl["config"],
            "solve_rate": solved / (float(Q) * world * K), "calls_in_flight": F,
            "budget": {"per_query_rounds": args.plan_rounds_per_query, "slots_per_call": min(args.plan_slots, Q),
                       "grow_calls_per_query": grow / (float(Q) * world * K),
                       "per_query_wall_s": args.plan_rounds_per_query * (gpu_ms / max(rounds / F, 1)) * 1e-3,
                       "note": "per-query budget (a query that spends it ends as Timeout and frees its slot); no per-batch deadline; "
                               "per_query_wall_s = budget x measured time per round, to set beside the reference arm's 1.0 s per query"},
            "e2e": {"value": solved / wall, "unit": PLAN_UNIT, "h2d_bytes_per_step": int(Q * 2 * desc.state_dim * 4 + Q * 4),
                    "d2h_bytes_per_step": int(Q * 8 + Q * kw["max_path"] * desc.state_dim * 4)},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "fp32", "scope": "all kernels of the planning rounds; FK + cull flops of the checked states only",
                         "achieved": achieved, "peak": peak_probe, "unit": "TFLOP/s", "frac": achieved / peak_probe if peak_probe else None,
                         "traffic": None, "states_checked_per_step": states / K / world, "rounds_per_step": rounds / K},
        }
        if not args.no_parity:  # every path returned by the first timed step re-validated by the oracle's motion validator
            from oracle import oracle as O

            pl = pls[0]
            status, paths = keep[warmup]
            a = [p[:-1] for p in paths if len(p) > 1]
            b = [p[1:] for p in paths if len(p) > 1]
            n_paths = len(a)
            bad = 0
            if n_paths:
                cnt = np.array([len(x) for x in a])
                valid, _ = O.check_motion_batch(desc, pl.lower_joint_limits, pl.upper_joint_limits, np.concatenate(a), np.concatenate(b),
                                                threads=os.cpu_count() or 1)
                owner = np.repeat(np.arange(n_paths), cnt)
                bad = int(len(set(owner[~valid].tolist())))
            starts, goals = sets[warmup % n_sets]
            ends_ok = sum(1 for i, p in enumerate(paths) if len(p) > 1 and np.allclose(p[0], starts[i], atol=1e-6) and np.allclose(p[-1], goals[i], atol=1e-6))
            rec["parity"] = {"paths_returned": n_paths, "paths_valid": n_paths - bad, "paths_start_goal_ok": int(ends_ok),
                             "oracle": "DiscreteMotionValidator restatement (oracle.c, float64, exhaustive leaf tests) over every edge of every path of one step"}
        if not args.no_cpu_baseline and world == 1:
            per = cpu_plans(wl, 1, 8, 5, 1)
            ns, dt = sum(x[0] for x in per), sum(x[1] for x in per)
            rec["cpu_baseline"] = {"value": ns / dt, "unit": PLAN_UNIT, "cores": 1, "kind": "port", "algorithm": "bvh",
                                   "solve_rate": ns / sum(x[2] for x in per), "grow_calls_per_query": sum(x[3] for x in per) / sum(x[2] for x in per),
                                   "sample": "8 queries of the same seeded workload one after the other on one thread, time limit 1 s per query"}
    for x in wls:
        c = x["world"]._cache
        if c is not None:
            c["dev"].close()
    torch.cuda.empty_cache()
    return rec


# ------------------------------------------------------------------------------------------------- reference arm
def reference_checks(name, args, batch, steps, warmup, budget_s):
    from mplib_b200 import workloads as WL
    from oracle import oracle as O

    wl = WL.get(name, batch)
    desc = wl["desc"]
    threads = os.cpu_count() or 1
    accel = O.Accel(desc)
    q = wl["make_q"](0, min(wl["batch"], 1 << 18)).astype(np.float64)
    per_step = []
    for s in range(warmup + steps):
        v, n, dt = cpu_leg(desc, q, threads, budget_s, accel)
        if s >= warmup:
            per_step.append((n, dt))
    n_tot, t_tot = sum(n for n, _ in per_step), sum(dt for _, dt in per_step)
    val = n_tot / t_tot
    return {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * t_tot / max(len(per_step), 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": wl["config"],
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "algorithm": "bvh",
                         "sample": f"{per_step[0][0] if per_step else 0} states/step of the same seeded workload; oracle.c hierarchical leg "
                                   f"(AABB-tree descent with early exit, edge-walking hull support), OpenMP x{threads}"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def reference_c4(args, batch, steps, warmup):
    from mplib_b200 import workloads as WL

    wl = WL.get("c4", batch)
    threads = os.cpu_count() or 1
    n_q = max(threads * 2, 16)
    per = cpu_plans(wl, threads, n_q, 5, warmup + steps)[warmup:]
    ns, dt = sum(x[0] for x in per), sum(x[1] for x in per)
    val = ns / dt
    return {
        "impl": "reference", "metric": PLAN_METRIC, "value": val, "unit": PLAN_UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * dt / max(len(per), 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": wl["config"],
        "solve_rate": ns / max(sum(x[2] for x in per), 1),
        "budget": {"per_query_time_s": 1.0, "grow_calls_per_query": sum(x[3] for x in per) / max(sum(x[2] for x in per), 1)},
        "cpu_baseline": {"value": val, "unit": PLAN_UNIT, "cores": threads, "kind": "port", "algorithm": "bvh",
                         "sample": f"{n_q} queries/step of the same seeded workload, one query per thread, OpenMP x{threads}, "
                                   "time limit 1 s per query"},
        "e2e": {"value": val, "unit": PLAN_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def also_list(args):
    if args.also.strip().lower() in ("", "none"):
        return []
    return [x for x in (y.strip() for y in args.also.split(",")) if x and x != args.workload]


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    head = args.workload
    if head == "c4":
        line = reference_c4(args, args.batch, args.steps, args.warmup)
    else:
        line = reference_checks(head, args, args.batch, args.steps, args.warmup, 1.0)
    also = {}
    for name in also_list(args):
        if name == "c4":
            also[name] = reference_c4(args, 0, 2, 1)
        else:
            also[name] = reference_checks(name, args, 0, 3, 1, 1.0)
    if also:
        line["also"] = also
    print(json.dumps(line))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    from mplib_b200 import _lib

    if _lib.load().mpb_device_count() == 0:
        raise SystemExit("bench.py: no CUDA device visible; the product path has no CPU fallback")
    ctx = Ctx()
    head = args.workload
    if head == "c4":
        line = run_c4(args, ctx, args.batch, args.steps, args.warmup)
    else:
        line = run_checks(head, args, ctx, args.batch, args.steps, args.warmup, True)
    also = {}
    for name in also_list(args):
        if name == "c4":
            also[name] = run_c4(args, ctx, 0, 4, 2)
        else:
            # C2 steps are 0.4 ms: 100 of them give a timed region long against the ranks' arrival skew and the one gather
            also[name] = run_checks(name, args, ctx, 0, 100 if name == "c2" else min(args.steps, 20), min(args.warmup, 3), False)
    if ctx.rank == 0:
        if also:
            line["also"] = also
        print(json.dumps(line))
    ctx.close()


if __name__ == "__main__":
    main()
