#!/usr/bin/env python
"""bench.py - particle-steps/s of the WCSPH particle-interaction path on the 3-D dam break.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--dp DP] [--real 8|4] [--impl reference]

A bench "step" is one advection step of the reference's main loop (tests/3d_examples/test_3d_dambreak/
dambreak.cpp:159-195): advection dt, density summation, the acoustic sub-step loop (Integration1stHalf +
Integration2ndHalf + AcousticTimeStep), then cell-list/sort and Inner+Contact relation updates.  The
metric is particle-steps/s = N_fluid * (#acoustic sub-steps) / time, everything else
amortised.  One JSON line on stdout (rank 0).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "particle-steps/sec, 3D dambreak WCSPH"
UNIT = "particle-steps/s"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_run(dp, steps, warmup):
    """the reference algorithm on the host cores: the CPU restatement (oracle, kind "port") - the TBB
    build of the reference cannot be compiled in this image (no Eigen/TBB/Boost/Simbody).  Runs on every
    core this process may use, whatever OMP_NUM_THREADS the launcher exported (torch.distributed.run sets 1)."""
    pk = importlib.import_module("sphinxsys-learning_b200")
    os.environ["OMP_NUM_THREADS"] = str(host_cores())
    from oracle import binding as oracle
    oracle.load().orc_set_num_threads(host_cores())
    c = pk.cases.DamBreak(oracle, ndim=3, dp=dp, real_bytes=8, with_observer=False)
    for _ in range(warmup):
        c.advection_step()
    n0 = c.n_acoustic
    t0 = time.perf_counter()
    for _ in range(steps):
        c.advection_step()
    dt = time.perf_counter() - t0
    sub = c.n_acoustic - n0
    return {"value": c.water.n * sub / dt, "unit": UNIT, "cores": int(oracle.load().orc_num_threads()), "kind": "port",
            "sample": f"3-D dam break dp={dp:g}: {c.water.n} fluid + {c.wall.n} wall particles, {steps} advection steps "
                      f"({sub} acoustic sub-steps) in {dt:.1f} s, OpenMP CPU restatement of the classic stored-neighbourhood algorithm"}, dt / max(steps, 1)


def slab_dp(dp1, world):
    """weak scaling: the same tank at a finer spacing, so every GPU keeps ~the single-GPU particle count
    (dp = 1/272 -> 20.1 M fluid particles on 1 GPU, 1/544 -> 161 M on 8; BASELINE.json configs[3])"""
    return 1.0 / round((1.0 / dp1) * world ** (1.0 / 3.0))


# stdout carries the one JSON line and nothing else: the process's fd 1 is pointed at stderr for the whole run
# (NCCL prints its version banner to stdout from C at the first communicator, whatever NCCL_DEBUG level is set)
_JSON_OUT = None


def claim_stdout():
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    claim_stdout()
    _JSON_OUT.write(line + "\n")
    _JSON_OUT.flush()


def slab_main(a, pk, stream, rank, world, local_rank, warmup):
    """N > 1: one x-slab per GPU with halo exchange + migration over NCCL; no replicas."""
    import torch
    import torch.distributed as dist
    from importlib import import_module
    slab = import_module("sphinxsys-learning_b200.slab")
    dp = slab_dp(a.dp, world)
    t_build = time.perf_counter()
    comm = slab.DistComm(torch.device("cuda", local_rank))
    case = slab.SlabDamBreak(pk.host, comm, dp=dp, real_bytes=a.real, device=local_rank, stream=stream.cuda_stream)
    t_build = time.perf_counter() - t_build
    P = case.parts[0]
    n_f = case.n_fluid()
    R = a.real

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sub0, l0 = case.n_acoustic, case.launches()
        t0 = time.perf_counter()
        ev0.record(stream)
        for _ in range(steps):
            fn()
        ev1.record(stream)
        barrier()
        ms = max(ev0.elapsed_time(ev1), 0.0)
        wall_ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([ms, wall_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0].item()), float(t[1].item()), case.n_acoustic - sub0, case.launches() - l0

    for _ in range(warmup):
        case.advection_step()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms_max, _, sub, launches = timed(case.advection_step, a.steps)
    clocks = sampler.stop() if sampler else None
    value = n_f * sub / (ms_max * 1e-3)

    # ---- roofline of the dominant kernel on this rank's slab (owned particles as i) ----
    w = P.case.water
    owned = P.oe - P.ob
    pairs = int(P.case.inner.neighbor_counts()[0][P.ob:P.oe].sum(dtype=np.int64)) + \
        int(P.case.wall_contact.neighbor_counts()[0][P.ob:P.oe].sum(dtype=np.int64))
    barrier()  # every rank times its kernel with the box otherwise quiet
    alg_bytes = owned * 20 * R + 4 * pairs
    reps = 5
    S = P.case.sys
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    S.integration_2nd_half(P.case.inner, P.case.wall_contact, True, 0.0)
    torch.cuda.synchronize()
    k0.record(stream)
    for _ in range(reps):
        S.integration_2nd_half(P.case.inner, P.case.wall_contact, True, 0.0)
    k1.record(stream)
    torch.cuda.synchronize()
    k_ms = k0.elapsed_time(k1) / reps
    peak, peak_src = peaks()
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": ("k_second_half_quad" if R == 8 else "k_second_half_tile") + " (Integration2ndHalf inner+wall) on rank 0's slab",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "kernel_ms": k_ms, "algorithmic_bytes_per_launch": alg_bytes,
                "pairs_per_particle": pairs / max(owned, 1)}
    a_step = 56 * R + 2 * 4 * pairs / max(owned, 1)
    step_frac = a_step * value / world / 1e9 / peak

    # ---- e2e: every step the slab's evolving state comes from pinned host arrays (double-buffered) ----
    e2e = None
    if not a.no_e2e:
        names = [nm for nm in slab.MIGRATE_VARS]
        n0 = w.n
        host = {nm: torch.from_numpy(w.get(nm)).pin_memory() for nm in names}
        h2d = sum(t.numel() * t.element_size() for t in host.values())

        def stage_inputs():
            for nm, ten in host.items():
                w.stage(nm, ten.data_ptr())

        def e2e_step():
            # this step's state was copied host -> device on the copy stream while the previous step computed; the host
            # rows (owned + halo of the captured arrangement) define the slab again, as a reload does: n back to the
            # captured count, commit (ids restart), the next upload starts, then cell list (owned pages) + relations +
            # halo maps and one advection step without its closing migration / re-decomposition (the next step's slab comes
            # from the host again and is configured then)
            w.resize(n0)
            stage_pending[0] or stage_inputs()
            w.commit_staged()
            stage_inputs()
            stage_pending[0] = True
            w.update_cell_list()
            P.case.update_configuration(observer=False)
            case._halo_maps()
            case.advection_step(reconfigure=False)
            case.mechanical_energy()

        stage_pending = [False]
        e2e_step()
        ms_e, wall_e, sub_e, _ = timed(e2e_step, a.steps)
        S.stage_wait()
        e2e = {"value": n_f * sub_e / (max(ms_e, wall_e) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": (8 + 8 * 6) * world,
               "note": "per rank: the slab's whole evolving fluid state (owned + halo rows, 11 variables) is uploaded from pinned host "
                       "arrays every step, double-buffered (sphb200_var_stage on a copy stream while the previous step computes, "
                       "sphb200_body_commit_staged with reload semantics), then cell list + relations + halo maps are rebuilt and one "
                       "advection step runs (Dt, density summation, the acoustic sub-steps with their halo refreshes; not the closing migration, "
                       "which the next reload replaces); read back: the all-reduced total mechanical "
                       "energy and the loop state; the host arrays hold the captured state and are not refreshed; bytes are rank 0's x N"}

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": warmup,
               "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f64" if R == 8 else "f32", "data": "synthetic",
               "config": {"workload": f"test_3d_dambreak lattice dp=1/{round(1 / dp)}: {n_f} fluid particles over {world} x-slabs "
                                      f"(rank 0: {owned} owned + {w.n - owned} halo fluid, {P.case.wall.n} wall)",
                          "acoustic_substeps": sub, "advection_steps": a.steps, "l2": "inputs larger than L2",
                          "parallelism": f"{world} slabs along x; per sub-step: halo records " + ("pushed into the neighbour GPU's memory by the pack kernel (NVLink peer stores, csrc/peer.cu)" if case.peer_halo else "over NCCL send/recv") + " and a device-side NCCL all-reduce of the dt maximum; per advection step: migration + fresh halo over NCCL, one sort",
                          "page_cuts": [int(c) for c in case.cuts], "build_seconds": round(t_build, 1)},
               "roofline": roofline, "step_roofline_frac": step_frac, "a_step_bytes": a_step,
               "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        emit(json.dumps(out))
    dist.destroy_process_group()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--dp", type=float, default=1.0 / 272.0, help="particle spacing (1/272 -> 20 123 648 fluid particles)")
    ap.add_argument("--real", type=int, default=8, choices=[4, 8], help="bytes per Real (8 = the reference's default build)")
    ap.add_argument("--impl", default="sphb200", choices=["sphb200", "reference"])
    ap.add_argument("--cpu-dp", type=float, default=1.0 / 96.0, help="spacing of the bounded CPU-baseline sample (1/96 -> 884 736 fluid particles)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-disordered", action="store_true", help="skip the jittered-state leg")
    ap.add_argument("--no-fp32", action="store_true", help="skip the fp32 build's legs (extra key \"fp32\")")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    warmup = max(a.warmup, 3)

    if a.impl == "reference":
        if rank != 0:
            return
        base, ms = cpu_reference_run(a.cpu_dp, max(a.steps, 1), a.warmup)
        emit(json.dumps({"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": a.gpus,
                          "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms * 1e3, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                          "config": {"workload": "test_3d_dambreak lattice, bounded CPU sample (host RAM bounds the stored-neighbourhood "
                                                 "algorithm to ~2 M particles; particle-steps/s of this O(N) algorithm does not depend on N)",
                                     "dp": a.cpu_dp, "cores": base["cores"]},
                          "cpu_baseline": base,
                          "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # keep stdout to the one JSON line: NCCL prints its version banner there at the VERSION/INFO levels
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pk = importlib.import_module("sphinxsys-learning_b200")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    if world > 1:
        return slab_main(a, pk, stream, rank, world, local_rank, warmup)

    out = single_gpu(a, pk, stream, local_rank, warmup, a.real, rank == 0 and not a.no_cpu_baseline)
    if not a.no_fp32 and a.real == 8:
        # the same workload and legs in the fp32 build (the reference's SPHINXSYS_USE_FLOAT), reported as an extra key
        f32 = single_gpu(a, pk, stream, local_rank, warmup, 4, False, short=True)
        out["fp32"] = {k: f32[k] for k in ("value", "ms_per_step", "roofline", "step_roofline_frac", "e2e", "value_disordered")}
    emit(json.dumps(out))


def single_gpu(a, pk, stream, device, warmup, R, with_cpu, short=False):
    import torch
    # ---- build the case ----
    t_build = time.perf_counter()
    case = pk.cases.DamBreak(pk.host, ndim=3, dp=a.dp, real_bytes=R, with_observer=True, device=device,
                             stream=stream.cuda_stream)
    S, w = case.sys, case.water
    t_build = time.perf_counter() - t_build
    n_f, n_w = w.n, case.wall.n
    steps = max(3, a.steps // 3) if short else a.steps
    evolving = ("Position", "Velocity", "Density", "Pressure", "DensityChangeRate", "Force", "ForcePrior", "Mass",
                "VolumetricMeasure", "PreviousGravityForce")

    def timed_steps(fn, k):
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0, sub0 = S.launches, case.n_acoustic
        t0 = time.perf_counter()
        ev0.record(stream)
        for _ in range(k):
            fn()
        ev1.record(stream)
        torch.cuda.synchronize()
        return ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3, case.n_acoustic - sub0, S.launches - l0

    def kernel_ms(fn, reps=5):
        """CUDA-event time of one call of fn on the launching stream; the fields fn accumulates into are restored"""
        keep = {nm: w.get(nm) for nm in ("DensityChangeRate", "Force", "Density", "Position", "Velocity", "Pressure")}
        fn()
        torch.cuda.synchronize()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record(stream)
        for _ in range(reps):
            fn()
        k1.record(stream)
        torch.cuda.synchronize()
        for nm, v in keep.items():
            w.set(nm, v)
        return k0.elapsed_time(k1) / reps

    # ---- `value`: state resident in HBM ---------------------------------------------------------
    for _ in range(warmup):
        case.advection_step()
    sampler = ClockSampler(device)
    ms, _, sub, launches = timed_steps(case.advection_step, steps)
    clocks = sampler.stop()
    value = n_f * sub / (ms * 1e-3)

    # ---- roofline of the dominant kernel (Integration2ndHalf: one launch per call) ---------------
    peak, peak_src = peaks()

    def roofline_now():
        pairs_in, pairs_w = case.inner.neighbor_counts()[1], case.wall_contact.neighbor_counts()[1]
        alg_bytes = n_f * 20 * R + 4 * (pairs_in + pairs_w)  # K3 of SURVEY 8d: 160 B fp64 / 80 B fp32 per particle + list
        k3 = kernel_ms(lambda: S.integration_2nd_half(case.inner, case.wall_contact, True, 0.0))
        k12 = kernel_ms(lambda: S.integration_1st_half(case.inner, case.wall_contact, True, 0.0))
        achieved = alg_bytes / (k3 * 1e-3) / 1e9
        st = case.inner.stats()
        return {"bound": "hbm", "kernel": "Integration2ndHalf inner+wall, fused init/update (" + ("k_second_half_quad: 2 lanes per group" if R == 8 else "k_second_half_tile") + "; wall sums from k_second_half_wall)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "kernel_ms": k3, "first_half_ms": k12, "algorithmic_bytes_per_launch": alg_bytes,
                "pairs_per_particle": (pairs_in + pairs_w) / n_f,
                "union_slots_per_group": st["union_entries"] / max(st["union_groups"], 1), "group_size": st["group_size"],
                "masked_search": st["masked_search"] and case.wall_contact.stats()["masked_search"]}

    roofline = roofline_now()
    try:  # DRAM bytes of this kernel at this size from the committed ncu --set full capture (profiles/traffic.json),
        # keyed by kernel name and N - a number from the capture, not from this run
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            tj = json.load(fh)
        key = f"k_second_half_quad<3,double,2,2>@{n_f}"
        if R == 8 and key in tj and os.environ.get("SPHB200_QUAD", "2") == "2":
            roofline["traffic"] = tj[key]["dram_bytes_per_launch"]
            roofline["traffic_source"] = "profiles/traffic.json (ncu --set full capture of this kernel at this N, not this run)"
    except Exception:
        pass
    a_step = 56 * R + 2 * 4 * roofline["pairs_per_particle"]  # SURVEY 8d A_step: 448/224 B fields + 2 list reads
    step_frac = a_step * value / 1e9 / peak

    # ---- `e2e`: host buffers in, observables out, every step ------------------------------------
    e2e = None
    if not a.no_e2e:
        host = {nm: torch.from_numpy(w.get(nm)).pin_memory() for nm in evolving}
        h2d = sum(v.numel() * v.element_size() for v in host.values())

        def stage_inputs():
            for nm, ten in host.items():
                w.stage(nm, ten.data_ptr())

        def e2e_step():
            # double-buffered inputs: this step's state was copied host -> device on the copy stream while the
            # previous step computed (the first one before the timed region); the next step's copy starts now.
            # Every timed step therefore issues one full upload, and the closing synchronisation waits for it.
            w.commit_staged()
            stage_inputs()
            # host arrays define the particle order: rebuild cell list + configuration as the
            # reference does after a reload (Dambreak.cpp:132-138)
            w.update_cell_list()
            case.update_configuration(observer=False)
            # the step without its closing re-configuration of the fluid relations: the next step's state comes from the
            # host again and is configured then; the probes are interpolated on the moved particles, which needs the
            # cell list of the new positions and the observer's contact relation only
            case.advection_step(reconfigure=False)
            w.update_cell_list()
            case.obs_contact.update()
            # the step's result: the observables the reference's regression tests record (energy, probe pressures)
            return case.mechanical_energy(), case.probe_pressure()

        stage_inputs()
        e2e_step()
        ms_e, wall_e, sub_e, _ = timed_steps(e2e_step, steps)
        # output-interval leg: what the reference moves when it writes a state (BodyStatesRecordingToVtp: Position,
        # Velocity, Density, Pressure) - asynchronous device -> pinned host copies, timed on their own
        out_names = ("Position", "Velocity", "Density", "Pressure")
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for nm in out_names:
            w.fetch_async(nm, host[nm].data_ptr())
        S.fetch_wait()
        out_ms = (time.perf_counter() - t0) * 1e3
        out_bytes = sum(host[nm].numel() * host[nm].element_size() for nm in out_names)
        e2e = {"value": n_f * sub_e / (max(ms_e, wall_e) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(8 + 8 * 6 + 8 * (sub_e // max(steps, 1) + 1)),
               "output_interval_leg": {"ms": out_ms, "d2h_bytes": int(out_bytes), "GBps": out_bytes / out_ms / 1e6,
                                       "what": "Position, Velocity, Density, Pressure to pinned host memory (sphb200_var_fetch_async + "
                                               "sphb200_ctx_fetch_wait), once per output interval in the reference, not per step"},
               "note": "every timed step uploads the whole evolving fluid state (10 variables) from pinned host arrays "
                       "(sphb200_var_stage + sphb200_body_commit_staged: double-buffered, the copy of step k+1's inputs runs on a "
                       "copy stream while step k computes; all copies lie inside the timed region, whose end waits for the device), "
                       "rebuilds cell list + relations, runs one advection step (Dt, density summation, the acoustic sub-steps), "
                       "re-bins the moved particles for the observer's contact relation and reads back the step's results: total "
                       "mechanical energy, the 6 probe pressures and the dt values.  The fluid relations are NOT rebuilt a second "
                       "time at the end of the step: the next step's state comes from the host and is configured then.  The host arrays hold the state captured "
                       "after the resident leg and are not refreshed, so every timed step integrates the same state; updated "
                       "particle state crosses back only at output intervals (output_interval_leg), as in the reference"}

    # ---- disordered state: SURVEY 8d jitter (U(-0.2 dp, 0.2 dp), seed 12345) + seeded velocities -------------------
    disordered = None
    if not a.no_disordered:
        rng = np.random.default_rng(12345)
        dtype = np.float64 if R == 8 else np.float32
        oid, _ = w.ids()
        pos = w.get("Position")
        # keyed by original id, so every build of the library sees the same state
        jit = (rng.uniform(-0.2, 0.2, size=pos.shape) * a.dp).astype(dtype)
        vel = rng.normal(0.0, 0.1 * case.U_ref, size=pos.shape).astype(dtype)
        w.set("Position", pos + jit[oid])
        w.set("Velocity", vel[oid])
        w.set("Density", np.full(n_f, case.rho0, dtype))
        for nm in ("DensityChangeRate", "Pressure"):
            w.set(nm, np.zeros(n_f, dtype))
        w.set("Force", np.zeros_like(pos))
        w.update_cell_list()
        case.update_configuration()
        case.dt = 0.0
        for _ in range(2):
            case.advection_step()
        ms_d, _, sub_d, _ = timed_steps(case.advection_step, steps)
        rd = roofline_now()
        disordered = {"value": n_f * sub_d / (ms_d * 1e-3), "ms_per_step": ms_d / steps, "acoustic_substeps": sub_d,
                      "kernel_ms": rd["kernel_ms"], "first_half_ms": rd["first_half_ms"], "frac": rd["frac"],
                      "pairs_per_particle": rd["pairs_per_particle"], "union_slots_per_group": rd["union_slots_per_group"],
                      "masked_search": rd["masked_search"],
                      "state": "lattice + U(-0.2 dp, 0.2 dp) jitter (seed 12345), velocities N(0, 0.1 U_ref), 2 untimed steps first"}

    cpu_base = None
    if with_cpu:
        cpu_base, _ = cpu_reference_run(a.cpu_dp, 16, 1)  # ~11 s of CPU work on 16 cores

    return {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if R == 8 else "f32", "data": "synthetic",
            "config": {"workload": f"test_3d_dambreak lattice dp={a.dp:.6g}: {n_f} fluid + {n_w} wall particles per GPU",
                       "acoustic_substeps": sub, "advection_steps": steps, "l2": "inputs larger than L2",
                       "parallelism": "single GPU", "build_seconds": round(t_build, 1)},
            "roofline": roofline, "step_roofline_frac": step_frac, "a_step_bytes": a_step,
            "value_disordered": disordered["value"] if disordered else None, "disordered": disordered,
            "cpu_baseline": cpu_base, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}


if __name__ == "__main__":
    main()
