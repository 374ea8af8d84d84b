#!/usr/bin/env python
"""bench.py -- agent-steps/s of the batched F1TENTH step (1080-beam lidar) on N B200s.

    python bench.py [--gpus N --steps K --warmup W] [--workload config3|config2|config4|config5] [--impl reference]

One "step" = one pass of the hot path (dynamics -> lidar -> iTTC -> [GJK, opponent ray-cast] -> laps) over
every env of the batch.  Workloads follow BASELINE.json `configs`:
  config3 (default; the one the metric is quoted on): 2^20 single-agent envs per GPU on example_map, fp32 lidar
           fast path, Philox noise, auto-reset from a raceline pose pool, random actions resident in HBM.
  config2: 4096 single-agent envs, fp64 parity path (numpy noise rows).
  config4: 65536 two-agent head-to-head envs, fp64 parity path.
Multi-GPU: pure env sharding, one process per GPU (torchrun), no collective inside step.  `value` is weak scaling
(per-GPU env count fixed); under torchrun the line also carries `strong_scaling` (BASELINE configs[2] as stated: 2^20
envs in total, N/G per GPU, through ShardedF110Env) and `collective` (the one NCCL call of the design: the all-gather of
episode statistics, after the timed region).  Timing: CUDA events on the launching stream, max over ranks.

`--impl reference` times the reference's own CPU path on the host cores: the UNMODIFIED F110Env (Numba/NumPy) from
oracle/_ref, one process per core (kind "reference"); if that cannot be imported on the box, the C restatement under
oracle/ with all host threads (kind "port"), and the line says why.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

WORKLOADS = {
    # name: (envs per GPU, agents, dtype, noise, auto_reset)
    "config3": (1 << 20, 1, "fp32", "philox", True),
    "config2": (4096, 1, "fp64", "numpy", False),
    "config4": (65536, 2, "fp64", "numpy", False),
}
L2_FLUSH_NOTE = "per-step working set (scans %.2f GB) exceeds the 126 MB L2; EDT (%.1f MB) stays L2-resident by design"


def load_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device_index):
        self.idx = device_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line)

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
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) != self.idx:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def make_poses(n_envs, agents):
    from red_gym_b200.host import raceline_poses
    lead = raceline_poses(n_envs)
    if agents == 1:
        return lead[:, None, :]
    rng = np.random.default_rng(4)
    out = [lead]
    for a in range(1, agents):
        lat = rng.uniform(-0.3, 0.3, n_envs)
        c = lead.copy()
        c[:, 0] += -1.5 * a * np.cos(lead[:, 2]) - lat * np.sin(lead[:, 2])
        c[:, 1] += -1.5 * a * np.sin(lead[:, 2]) + lat * np.cos(lead[:, 2])
        out.append(c)
    return np.stack(out, axis=1)


def action_bank(torch, n_envs, agents, device, n_bank, seed):
    """Synthetic action stream: steer ~ U(-0.4189, 0.4189), speed ~ U(0, 8); a bank of tensors
    generated before the timed region so that inputs are resident in HBM when it starts."""
    g = torch.Generator(device=device).manual_seed(seed)
    bank = []
    for _ in range(n_bank):
        a = torch.rand((n_envs, agents, 2), generator=g, device=device, dtype=torch.float64)
        a[..., 0] = (a[..., 0] * 2 - 1) * 0.4189
        a[..., 1] = a[..., 1] * 8.0
        bank.append(a.contiguous())
    return bank


def l2_probe(device_index):
    """Measured L2 read bandwidth of this GPU (f110_l2_probe: hand-written read-only kernels, L1 bypassed, CUDA events):
    whole 128-byte lines and random 4-byte gathers (one 32-byte sector per lane), over a 10 MB table (the fp32 EDT of
    example_map) and a 26 MB one (its padded copy, what k_raycast32 reads).  Both stay L2-resident (126 MB)."""
    import ctypes
    from red_gym_b200 import _native as nat
    out = {}
    for mb in (10, 26):
        r = (ctypes.c_double * 6)()
        nat.check(nat.load().f110_l2_probe(device_index, mb << 20, 5, r, None))
        out["%dMB" % mb] = {"stream_GBps": r[0], "gather_sectors_per_s": r[1], "gather_sector_GBps": r[1] * 32 / 1e9,
                            "gather_useful_GBps": r[1] * 4 / 1e9, "stream_ms": r[2], "gather_ms": r[3]}
    out["how"] = "f110_l2_probe: ld.global.cg (L1 bypassed), 8 independent loads per thread, 32 (stream) / 64 (gather) warps per SM, 5 launches each"
    return out


def pcie_d2h_peak(torch, dist, device, world, dst, chunks=8, reps=3):
    """The roof of the e2e figure: plain device -> page-locked host copies into `dst` (the pinned scan buffer of the e2e leg,
    same bytes), one cudaMemcpyAsync per chunk, no kernels running; under torchrun every rank copies at the same time (max
    over ranks)."""
    dst = dst.reshape(-1)
    n, nbytes = dst.numel(), dst.numel() * dst.element_size()
    src = torch.empty(n, dtype=dst.dtype, device=device)
    per = (n + chunks - 1) // chunks
    def once():
        for c in range(chunks):
            dst[c * per:(c + 1) * per].copy_(src[c * per:(c + 1) * per], non_blocking=True)
        torch.cuda.synchronize()
    once()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    t = torch.tensor([time.perf_counter() - t0], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del src
    gbps = reps * nbytes * world / float(t.item()) / 1e9
    return {"GBps_all_ranks": gbps, "GBps_per_rank": gbps / world, "bytes_per_rank": nbytes, "chunks": chunks, "ranks_copying_at_once": world}


def kernel_profile(dtype):
    """What one ncu --set full capture of the bench-size launch measured for the ray-cast kernel (profiles/roofline_traffic.json,
    written by tools/ncu_roofline.py from the committed .csv): per-scan DRAM bytes, L2 read sectors, warp instructions, issue
    activity.  ncu counters cannot be read live; everything time-like in the bench line is measured live with CUDA events."""
    tp = os.path.join(REPO, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        return json.load(open(tp)).get(dtype)
    return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    from red_gym_b200 import BatchF110Env, MAPS_DIR

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa_cpus = None
    if world > 1:      # one process per GPU: host buffers on the GPU's own NUMA node
        from red_gym_b200.host import bind_to_gpu_numa
        numa_cpus = bind_to_gpu_numa(local)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    n_envs, agents, dtype, noise, auto = WORKLOADS[args.workload]
    if args.envs:
        n_envs = args.envs
    if args.dtype:
        dtype = args.dtype
        noise = "numpy" if dtype == "fp64" else "philox"
    B = 1080
    env_kw = dict(device=local, dtype=dtype, noise=noise, num_agents=agents, auto_reset=True, reset_poses=make_poses(1024, agents),
                  raycast_variant=args.variant, map=os.path.join(MAPS_DIR, "example_map"), seed=12345)
    env = BatchF110Env(num_envs=n_envs, env_offset=rank * n_envs, **env_kw)
    # every workload auto-resets here: a throughput run must keep cars alive (the parity tests cover no-reset runs)
    env.reset(make_poses(n_envs, agents))
    bank = action_bank(torch, n_envs, agents, device, 8, 2024 + rank)
    K, W = args.steps, args.warmup

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(e, actions, k):
        """k steps of env e bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks (ms)."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(k):
            e.step(actions[i % len(actions)])
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # SURVEY 8d config 3 asks for 200 warm-up steps: right after a reset the cars stand on their start grid, in track
    # order and slow, and a step is ~5 % cheaper than in the steady state (env order decorrelated, cars at speed,
    # auto-resets under way).  Whatever W is, the timed region starts in the steady state.
    preroll = max(0, 200 - W) if args.preroll < 0 else args.preroll
    for i in range(preroll + W):
        env.step(bank[i % len(bank)])
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = env.launch_count
    ms_max = timed_steps(env, bank, K)
    launches = env.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    total_agent_steps = float(n_envs) * agents * K * world
    value = total_agent_steps / (ms_max * 1e-3)

    # ---- roofline of the dominant kernel (ray-cast), measured live with CUDA events on the launching stream
    env.enable_timing(True)
    rc_ms = []
    for i in range(max(3, min(K, 10))):
        env.step(bank[i % len(bank)])
        rc_ms.append(env.last_raycast_ms())
    env.enable_timing(False)
    env.count_lookups(True)
    n_cnt = 3
    for i in range(n_cnt):
        env.step(bank[i % len(bank)])
    torch.cuda.synchronize()
    kernel_lookups_per_scan = env.lookups() / float(n_cnt * n_envs * agents)
    # reference definition: every distance_transform call counts, including the first one of each ray;
    # the kernel fetches the scan origin once per scan instead of once per beam
    lookups_per_scan = kernel_lookups_per_scan + (B - 1)
    env.count_lookups(False)
    esz = 4 if dtype == "fp32" else 8
    scans_per_launch = n_envs * agents
    bytes_per_scan = lookups_per_scan * esz + B * esz            # L*sizeof(edt_t) + B*sizeof(scan_t)
    rc_avg_ms = float(np.mean(rc_ms))
    rc_s = rc_avg_ms * 1e-3
    achieved = bytes_per_scan * scans_per_launch / rc_s / 1e9
    hbm_peak, hbm_src = load_peaks()
    l2 = l2_probe(local) if rank == 0 else None
    roofline = None
    if rank == 0:
        tab = "26MB" if dtype == "fp32" else "10MB"       # the table the kernel gathers from: padded fp32 EDT 26 MB, fp64 EDT 20.5 MB
        l2_peak = l2[tab]["stream_GBps"]
        prof = kernel_profile(dtype)
        roofline = {"kernel": "k_raycast32" if dtype == "fp32" else "k_raycast64", "bound": "l2", "achieved": achieved, "peak": l2_peak,
                    "unit": "GB/s", "frac": achieved / l2_peak,
                    "peak_source": "measured live: f110_l2_probe, 128-byte-line reads of an L2-resident %s table" % tab,
                    "traffic": prof["dram_bytes_per_scan"] * scans_per_launch if prof else None,
                    "launch_ms": rc_avg_ms, "lookups_per_scan": lookups_per_scan, "kernel_lookups_per_scan": kernel_lookups_per_scan,
                    "bytes_per_scan": bytes_per_scan, "scans_per_launch": scans_per_launch, "raycast_share_of_step": rc_avg_ms / (ms_max / K),
                    "lookups_per_s": kernel_lookups_per_scan * scans_per_launch / rc_s,
                    "note": "algorithmic bytes (SURVEY 8d) over the measured L2 read peak.  The kernel is NOT bound by L2 or HBM bandwidth: "
                            "L1 serves most gathers (l2_sectors / hbm below are what ncu measured leaving L1 / L2); it is bound by the latency "
                            "of its dependent L1-hit gather chain and by issue slots (issue below)"}
        if prof and "l2_read_sectors_per_scan" in prof:
            sect = prof["l2_read_sectors_per_scan"] * scans_per_launch
            roofline["l2_sectors"] = {"read_GBps": sect * 32 / rc_s / 1e9, "gather_peak_GBps": l2[tab]["gather_sector_GBps"],
                                      "frac_of_gather_peak": sect / rc_s / l2[tab]["gather_sectors_per_s"],
                                      "l1_hit_pct": prof.get("l1_hit_pct"), "source": prof.get("source")}
            roofline["hbm"] = {"traffic_GBps": prof["dram_bytes_per_scan"] * scans_per_launch / rc_s / 1e9, "peak": hbm_peak,
                               "frac": prof["dram_bytes_per_scan"] * scans_per_launch / rc_s / 1e9 / hbm_peak, "peak_source": hbm_src}
            # instruction-limited floor of this algorithm: the 10-instruction marching loop at 32 of 32 lanes and nothing else
            floor = kernel_lookups_per_scan * 10.0 / 32.0
            roofline["issue"] = {"warp_inst_per_scan": prof["warp_inst_per_scan"], "issue_active_pct": prof["issue_active_pct"],
                                 "loop_only_warp_inst_per_scan": floor, "frac": floor / prof["warp_inst_per_scan"] * prof["issue_active_pct"] / 100.0,
                                 "thread_inst_per_lookup": prof.get("thread_inst_per_lookup"), "loop_lane_efficiency": prof.get("loop_lane_efficiency")}

    # ---- launch-bound batches: the same steps replayed from a CUDA graph (informational; `value` stays the eager API figure)
    graph_replay = None
    if n_envs * agents <= 65536 and not args.no_graph:
        G = 8
        gacts = torch.stack(bank[:G])
        env.reserve_noise(G * (K // G + 8) + 16)
        cg = env.capture_step_graph(gacts, steps=G, warmup=3)
        reps = max(1, K // G)
        for _ in range(2):
            cg.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(reps):
            cg.replay()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        env.note_replayed(G * (reps + 2))
        graph_replay = {"value": float(n_envs) * agents * G * reps * world / (float(t.item()) * 1e-3), "unit": "agent-steps/s",
                        "ms_per_step": float(t.item()) / (G * reps), "steps_per_graph": G, "replays": reps,
                        "api": "BatchF110Env.capture_step_graph (torch.cuda.CUDAGraph of f110_step launches)"}

    # ---- e2e: the same step through f110_step_host with HOST buffers (H2D actions, D2H observation every step)
    e2e = None
    if not args.no_e2e:
        sdt = torch.float32 if dtype == "fp32" else torch.float64
        act_h = torch.zeros((n_envs, agents, 2), dtype=torch.float64).pin_memory()
        obs_h = torch.zeros((10, n_envs, agents), dtype=torch.float64).pin_memory()
        scans_h = torch.zeros((n_envs, agents, B), dtype=sdt).pin_memory()
        done_h = torch.zeros((n_envs,), dtype=torch.uint8).pin_memory()
        act_h.copy_(bank[0])
        Ke = max(2, min(K, args.e2e_steps))

        def host_steps(e, scans_buf):
            for _ in range(2):
                e.step_host(act_h, obs_h, scans_buf, done_h)
            barrier()
            t0 = time.perf_counter()
            for _ in range(Ke):
                e.step_host(act_h, obs_h, scans_buf, done_h)
            torch.cuda.synchronize()
            t = torch.tensor([time.perf_counter() - t0], device=device, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(n_envs) * agents * Ke * world / float(t.item())

        d2h = (obs_h.numel() * 8 + scans_h.numel() * scans_h.element_size() + done_h.numel())
        e2e = {"value": host_steps(env, scans_h), "unit": "agent-steps/s", "h2d_bytes_per_step": act_h.numel() * 8 * world,
               "d2h_bytes_per_step": d2h * world, "steps": Ke,
               "api": "f110_step_host (C ABI, pinned host buffers, full observation incl. scans as %s copied back every step)" % ("float32" if dtype == "fp32" else "float64"),
               "host_numa_binding": "rank pinned to %d GPU-local CPUs" % len(numa_cpus) if numa_cpus else "none"}
        e2e["d2h_GBps"] = e2e["value"] / (n_envs * agents * world) * d2h * world / 1e9
        # informational: the same call with scans_host = NULL (scans stay in HBM for a device-side consumer such as the
        # bitmap rasteriser / a policy; poses, velocities, flags and done still come back every step)
        e2e["scans_on_device"] = {"value": host_steps(env, None), "unit": "agent-steps/s", "d2h_bytes_per_step": (obs_h.numel() * 8 + done_h.numel()) * world}
        # the roof of the full-observation figure: the same bytes over the same link(s) with nothing else going on
        roof = pcie_d2h_peak(torch, dist, device, world, scans_h)
        e2e["pcie_d2h_peak"] = roof
        e2e["frac_of_pcie_peak"] = e2e["d2h_GBps"] / roof["GBps_all_ranks"]
        # opt-in compact observation (F110_SCAN_U16: ranges as uint16 at 0.5 mm per LSB, converted in the ray-cast epilogue):
        # half the bytes over PCIe.  NOT the headline: the observation differs from the reference's by <= 0.25 mm per beam.
        if dtype == "fp32" and agents == 1 and not args.no_compact:
            del env
            torch.cuda.empty_cache()
            cenv = BatchF110Env(num_envs=n_envs, env_offset=rank * n_envs, scan_format="u16", **env_kw)
            cenv.reset(make_poses(n_envs, agents))
            for i in range(20):
                cenv.step(bank[i % len(bank)])
            u16 = scans_h.view(torch.uint16).reshape(-1)[:n_envs * agents * B].reshape(n_envs, agents, B)      # the same pinned pages
            cd2h = obs_h.numel() * 8 + u16.numel() * 2 + done_h.numel()
            cv = host_steps(cenv, u16)
            e2e["compact_u16"] = {"value": cv, "unit": "agent-steps/s", "d2h_bytes_per_step": cd2h * world,
                                  "d2h_GBps": cv / (n_envs * agents * world) * cd2h * world / 1e9,
                                  "api": "f110_step_host on a handle created with scan_format = F110_SCAN_U16",
                                  "tolerance": "|range - fp32 range| <= 0.25 mm (0.5 mm per LSB), negative noisy ranges stored as 0; states and flags bit-identical (tests/test_gpu_abi_v2.py)",
                                  "device_ms_per_step": timed_steps(cenv, bank, max(3, min(K, 10))) / max(3, min(K, 10))}
            cenv.close()
            del cenv
        del scans_h, obs_h

    # ---- multi-GPU as BASELINE configs[2] states it: 2^20 envs in total, N/G per GPU (strong scaling), and the design's
    # one collective: the all-gather of episode statistics over NCCL (never inside step)
    strong, collective = None, None
    if world > 1 and args.workload == "config3":
        from red_gym_b200.sharding import ShardedF110Env
        try:
            del env
        except NameError:
            pass
        torch.cuda.empty_cache()
        n_total = args.envs or WORKLOADS[args.workload][0]
        kw = dict(env_kw)
        kw.pop("device")
        sh = ShardedF110Env(n_total, rank=rank, world=world, device=local, **kw)
        sh.reset(make_poses(n_total, agents))
        sbank = action_bank(torch, sh.count, agents, device, 8, 4048 + rank)
        for i in range(preroll + W):
            sh.step(sbank[i % len(sbank)])
        s_ms = timed_steps(sh, sbank, K)
        strong = {"value": float(n_total) * agents * K / (s_ms * 1e-3), "unit": "agent-steps/s", "scaling": "strong", "envs_total": n_total,
                  "envs_per_gpu": sh.count, "ms_per_step": s_ms / K, "api": "red_gym_b200.sharding.ShardedF110Env",
                  "limiter": "per-scan cost of the ray-cast at the smaller per-GPU batch (DESIGN.md 9); no collective in step"}
        # gather_stats(): every rank receives the (n_total, A, 5) statistics of all envs in global env order
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = sh.gather_stats()        # first call sets up the NCCL communicator / buffers
        barrier()
        e0.record()
        st = sh.gather_stats()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # global order check: column 4 (episodes) and the others must equal what each owner holds; rank 0 compares the slice
        # of every rank against that rank's own copy, fetched with a broadcast
        from red_gym_b200.sharding import episode_stats, shard_range
        ok = True
        for r in range(world):
            b_, c_ = shard_range(n_total, r, world)
            own = episode_stats(sh.env) if r == rank else torch.empty((c_, agents, 5), dtype=torch.float64, device=device)
            dist.broadcast(own, src=r)
            ok = ok and bool(torch.equal(st[b_:b_ + c_], own))
        collective = {"op": "all_gather_into_tensor (NCCL)", "what": "episode statistics of all envs (lap_count, lap_time, collided, sim_time, episodes)",
                      "bytes_per_rank_out": int(st.numel() * 8), "ms": float(t.item()), "global_order_verified": ok, "ranks": world,
                      "algbw_GBps": st.numel() * 8 / (float(t.item()) * 1e-3) / 1e9}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_baseline = cpu_reference(args.workload, seconds=12.0)
    if rank == 0:
        scan_gb = n_envs * agents * B * esz / 1e9
        line = {"metric": "agent-steps/sec (1080-beam lidar)", "value": value, "unit": "agent-steps/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32" if dtype == "fp32" else "f64", "data": "synthetic",
                "config": {"workload": args.workload, "envs_per_gpu": n_envs, "agents": agents, "beams": B, "map": "example_map",
                           "integrator": "RK4", "noise": noise, "auto_reset": True, "parallelism": "env-shard x%d" % world,
                           "raycast_variant": args.variant, "untimed_steps_before_timing": preroll + W,
                           "l2_flush": L2_FLUSH_NOTE % (scan_gb, 10.24 if dtype == "fp32" else 20.48)},
                "roofline": roofline, "l2_roofline": l2, "cpu_baseline": cpu_baseline, "e2e": e2e, "graph_replay": graph_replay, "strong_scaling": strong,
                "collective": collective, "gpu_launches": int(launches), "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_sweep(args):
    """BASELINE config 5: ray-cast only, B in {270, 1080, 2160} x EDT size {1024, 2048, 4096}^2 (example_map resampled),
    2^18 poses per GPU uniform over free space (clearance 0.3 m), through f110_raycast.  One JSON line, `sweep` = list
    of points (scans/s, rays/s, EDT lookups/s, algorithmic GB/s); L2 hit rate and DRAM bytes per point come from the
    ncu captures under profiles/."""
    import torch
    import torch.distributed as dist
    from red_gym_b200 import BatchF110Env
    from red_gym_b200 import host
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    n = args.envs or (1 << 18)
    sizes = [int(x) for x in args.sweep_sizes.split(",")]
    beams = [int(x) for x in args.sweep_beams.split(",")]
    dtypes = [args.dtype] if args.dtype else ["fp32", "fp64"]
    K, W = max(args.steps, 1), max(args.warmup, 3)
    points = []
    for size in sizes:
        m = host.resampled_map(size)
        poses = torch.from_numpy(host.free_space_poses(m, n, seed=7 + rank)).to(device)
        for dtype in dtypes:
            esz = 4 if dtype == "fp32" else 8
            for B in beams:
                env = BatchF110Env(num_envs=1, num_agents=1, device=local, dtype=dtype, noise="off", num_beams=B, map_data=m,
                                   raycast_variant=args.variant)
                out = torch.empty((n, B), dtype=torch.float32 if dtype == "fp32" else torch.float64, device=device)
                for _ in range(W):
                    env.raycast(poses, out)
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(K):
                    env.raycast(poses, out)
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / K], device=device, dtype=torch.float64)
                if world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
                env.count_lookups(True)
                env.raycast(poses, out)
                torch.cuda.synchronize()
                lk = env.lookups() / float(n) + (B - 1)          # reference definition: one lookup per distance_transform call
                env.count_lookups(False)
                scans_s = n * world / (ms * 1e-3)
                points.append({"edt": size, "beams": B, "dtype": "f32" if dtype == "fp32" else "f64", "ms_per_launch": ms,
                               "scans_per_s": scans_s, "rays_per_s": scans_s * B, "lookups_per_scan": lk,
                               "lookups_per_s": scans_s * lk, "algorithmic_GBps": scans_s * (lk + B) * esz / 1e9,
                               "edt_MB": size * size * esz / 1e6})
                env.close()
                del out
    if rank == 0:
        peak, peak_src = load_peaks()
        l2 = l2_probe(local)
        for p in points:      # per GPU: algorithmic GB/s against the measured L2 read peak (26 MB table; both tables stay L2-resident)
            p["algorithmic_GBps_per_gpu"] = p["algorithmic_GBps"] / world
            p["frac_of_l2_stream_peak"] = p["algorithmic_GBps_per_gpu"] / l2["26MB"]["stream_GBps"]
            p["lookups_vs_l2_gather_peak"] = p["lookups_per_s"] / world / l2["26MB"]["gather_sectors_per_s"]
        print(json.dumps({"metric": "scans/s (ray-cast only sweep, config 5)", "unit": "scans/s", "n_gpus": world, "steps": K, "warmup": W,
                          "value": max(p["scans_per_s"] for p in points), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "data": "synthetic", "config": {"workload": "config5", "poses_per_gpu": n, "map": "example_map resampled",
                                                          "pose_sampling": "uniform over free cells with clearance 0.3 m, seed 7 + rank"},
                          "hbm_peak_GBps": peak, "peak_source": peak_src, "l2_roofline": l2,
                          "note": "lookups_vs_l2_gather_peak > 1 means L1 serves part of the gathers; L2 hit rate / sectors / DRAM bytes per point: ncu pass under profiles/",
                          "sweep": points}))
    if world > 1:
        dist.destroy_process_group()


def run_downstream(args):
    """The two device-side neighbours of the step: scan -> 256x256 bitmap (FILL, the call src/SAL.py makes
    every step) and the pure-pursuit planner, timed on the scans / poses of a stepped env.  One JSON line."""
    import torch
    from red_gym_b200 import BatchF110Env, BatchPurePursuitPlanner, LidarBitmap, MAPS_DIR
    from argparse import Namespace
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    n = args.envs or 65536
    K, W = max(args.steps, 1), max(args.warmup, 3)
    env = BatchF110Env(num_envs=n, device=local, dtype="fp32", noise="philox", num_agents=1, map=os.path.join(MAPS_DIR, "example_map"))
    env.reset(make_poses(n, 1))
    bank = action_bank(torch, n, 1, device, 2, 2024)
    for i in range(20):
        env.step(bank[i % 2])
    scans = env.scans.view(n, 1080)
    peak, peak_src = load_peaks()
    res = {}

    def timeit(fn):
        for _ in range(W):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / K

    bm = LidarBitmap(num_beams=1080, draw_mode="FILL", bg_color="black", scan_dtype="fp32", device=local)
    out = torch.empty((n, 256, 256), dtype=torch.uint8, device=device)
    ms = timeit(lambda: bm(scans, out=out))
    by = n * (256 * 256 + 600 * 4)            # algorithmic bytes: the image written + the 600 sampled ranges read
    res["bitmap_u8"] = {"ms": ms, "images_per_s": n / (ms * 1e-3), "algorithmic_GBps": by / (ms * 1e-3) / 1e9,
                        "frac_of_hbm_peak": by / (ms * 1e-3) / 1e9 / peak, "bytes_per_image": by // n}
    outp = torch.empty((n, 256, 8), dtype=torch.int32, device=device)
    ms = timeit(lambda: bm(scans, out=outp, packed=True))
    by = n * (256 * 32 + 600 * 4)
    res["bitmap_packed"] = {"ms": ms, "images_per_s": n / (ms * 1e-3), "algorithmic_GBps": by / (ms * 1e-3) / 1e9,
                            "frac_of_hbm_peak": by / (ms * 1e-3) / 1e9 / peak, "bytes_per_image": by // n}
    conf = Namespace(wpt_path=os.path.join(MAPS_DIR, "example_waypoints.csv"), wpt_delim=";", wpt_rowskip=3, wpt_xind=1, wpt_yind=2, wpt_vind=5)
    pl = BatchPurePursuitPlanner(conf, 0.17145 + 0.15875, device=local)
    act = torch.empty((n, 1, 2), dtype=torch.float64, device=device)
    ms = timeit(lambda: pl.plan_actions(env, 0.82461887897713965, 1.375, out=act))
    res["pure_pursuit"] = {"ms": ms, "plans_per_s": n / (ms * 1e-3), "segments_per_plan": 782}
    # the same poses 16 times over in one launch: at 65 536 cars the call is shorter than its own host-side overhead
    big = [env.state[k].reshape(-1).repeat(16).contiguous() for k in (0, 1, 4)]
    ms = timeit(lambda: pl.plan(big[0], big[1], big[2], 0.82461887897713965, 1.375))
    res["pure_pursuit_16x"] = {"ms": ms, "plans": 16 * n, "plans_per_s": 16 * n / (ms * 1e-3)}
    cpu = None
    if not args.no_cpu:
        sys.path.insert(0, os.path.join(REPO, "oracle"))
        import bitmap_oracle as bo
        import pursuit_oracle as po
        sc = scans[:64].cpu().numpy()
        t0 = time.perf_counter(); k = 0
        while time.perf_counter() - t0 < 8.0 and k < 64:
            bo.lidar_to_bitmap(sc[k], draw_mode="FILL", bg_color="black"); k += 1
        tb = (time.perf_counter() - t0) / k
        wp = po.load_waypoints(conf.wpt_path)
        st = env.state.cpu().numpy()
        t0 = time.perf_counter(); j = 0
        while time.perf_counter() - t0 < 4.0 and j < 2000:
            po.plan(wp, st[0, j, 0], st[1, j, 0], st[4, j, 0], 0.82461887897713965, 1.375, 0.3302); j += 1
        tp = (time.perf_counter() - t0) / j
        cpu = {"kind": "port", "cores": 1, "bitmap_images_per_s": 1.0 / tb, "pursuit_plans_per_s": 1.0 / tp,
               "sample": "%d images / %d plans with the pure-Python + numpy oracle (the reference uses OpenCV / Numba)" % (k, j)}
    print(json.dumps({"metric": "images/s (scan -> 256x256 bitmap, FILL)", "value": res["bitmap_u8"]["images_per_s"], "unit": "images/s",
                      "n_gpus": 1, "steps": K, "warmup": W, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                      "data": "synthetic", "config": {"workload": "downstream", "scans": n, "beams": 1080, "target_beam_count": 600},
                      "hbm_peak_GBps": peak, "peak_source": peak_src, "downstream": res, "cpu_baseline": cpu}))


def cpu_port(workload, seconds=12.0, nthreads=0):
    """The C restatement of the reference (oracle/f110_oracle.c, pthreads over envs) on a bounded sample of the workload.
    Imports nothing of the product: the maps are found by path."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import f110_oracle as fo
    import ref_runner
    maps_dir = os.path.join(REPO, "red_gym_b200", "maps")
    _, agents, _, _, _ = WORKLOADS[workload]
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    n = 64 * cores
    poses = ref_runner._poses(os.path.join(maps_dir, "example_waypoints.csv"), n, agents)
    env = fo.OracleEnv(n, agents, map=os.path.join(maps_dir, "example_map"), nthreads=nthreads or cores, noise_rows=4096)
    env.reset(poses)
    rng = np.random.default_rng(2024)

    def act():
        return np.stack([rng.uniform(-0.4189, 0.4189, (n, agents)), rng.uniform(0, 8, (n, agents))], axis=-1)
    for _ in range(3):
        env.step(act())
    steps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        env.step(act())
        steps += 1
        d = env.done.astype(bool)
        if d.any():     # auto-reset like the GPU workload
            env.reset(poses, mask=d)
    dt = time.perf_counter() - t0
    return {"value": n * agents * steps / dt, "unit": "agent-steps/s", "cores": nthreads or cores, "kind": "port",
            "cpu_model": ref_runner.cpu_model(), "host_cpus": os.cpu_count(), "timed_seconds": dt,
            "sample_envs": n, "sample_ms_per_step": 1e3 * dt / max(steps, 1),
            "sample": "C restatement of the reference (oracle/f110_oracle.c): %d envs x %d steps (%.1f s) of %s on example_map, fp64, numpy noise rows, host auto-reset" % (n, steps, dt, workload)}


def cpu_reference(workload, seconds=12.0):
    """The reference's own CPU implementation of the path on the box's host cores: the unmodified F110Env, one process per
    core (kind "reference"); falls back to the C port (kind "port") when the reference cannot run here, saying why."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import ref_runner
    _, agents, _, _, _ = WORKLOADS[workload]
    ok, why = ref_runner.available()
    if ok:
        try:
            cb = ref_runner.run(agents=agents, seconds=seconds, solo_seconds=4.0)
            cb["sample_ms_per_step"] = 1e3 * cb["timed_seconds"] * cb["cores"] / max(cb["sample_steps"], 1)      # one step of every process
            cb["sample_envs"] = cb["cores"]
            return cb
        except Exception as e:
            why = "reference run failed: %s" % e
    cb = cpu_port(workload, seconds=seconds)
    cb["why_not_reference"] = why
    return cb


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    seconds = max(5.0, min(60.0, 1.5 * (args.steps + args.warmup)))
    cb = cpu_reference(args.workload, seconds=seconds)
    n_envs, agents, dtype, noise, auto = WORKLOADS[args.workload]
    line = {"impl": "reference", "metric": "agent-steps/sec (1080-beam lidar)", "value": cb["value"], "unit": "agent-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["sample_ms_per_step"],
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "envs_per_step": cb["sample_envs"], "agents": agents, "beams": 1080, "map": "example_map",
                       "integrator": "RK4", "noise": "numpy", "auto_reset": True,
                       "note": "steps / warmup are echoed from the command line; this arm runs for a fixed time instead: %.1f s timed (%.0f s requested "
                               "= 1.5 x (steps + warmup), clamped to 5..60) after 200 warm-up steps per process; a step of this arm = one step of every "
                               "process's env (one F110Env per host core)" % (cb["timed_seconds"], seconds)},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS) + ["config5", "downstream"])
    ap.add_argument("--sweep-sizes", default="1024,2048,4096")
    ap.add_argument("--sweep-beams", default="270,1080,2160")
    ap.add_argument("--envs", type=int, default=0, help="override envs per GPU")
    ap.add_argument("--dtype", default=None, choices=["fp32", "fp64"])
    ap.add_argument("--variant", type=int, default=0, help="raycast kernel variant bits (64: no spatial sort, 128: no CTA-per-scan split, ...)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--preroll", type=int, default=-1, help="untimed steps before the W warm-up steps (default: up to 200 in total)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-compact", action="store_true", help="skip the uint16 compact-observation e2e leg")
    ap.add_argument("--no-graph", action="store_true", help="skip the CUDA-graph replay leg of small workloads")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.workload == "config5":
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "the sweep is a GPU-only diagnostic; use --workload config3"}))
        else:
            run_sweep(args)
    elif args.workload == "downstream":
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "diagnostic workload; use --workload config3"}))
        else:
            run_downstream(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
