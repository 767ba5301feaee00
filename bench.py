#!/usr/bin/env python
"""bench.py -- 1080p open-loop motion-estimation throughput (BASELINE.json configs[1]).

Workload (one "step"): the ME stage of a 121-picture 1920x1080 8-bit sequence coded with the
reference's hierarchical mini-GOP 8 / 4 temporal layers prediction structure
(EbPredictionStructure.c:183-240): picture 0 is intra (pyramid only), pictures 1..120 are B pictures
-> 121 pyramid builds (PictureAnalysis decimation + padding) and 120 x MotionEstimateLcu-per-SB
(510 SBs x 2 lists x 85 PUs each). Synthetic, seeded luma.

  value  frames/s with all luma planes already resident in HBM (results stay in HBM)
  e2e    frames/s through the C-ABI with HOST buffers: H2D of every luma plane from pinned memory and
         D2H of every picture's me_results image inside the timed region
  roofline  of the dominant kernel (integer full-pel search): INT-pipe bound, see DESIGN.md
  cpu_baseline  the reference's own AVX2 ME (oracle/_ref, compiled from /root/reference) timed on the
         host cores of this box on a bounded sample of the same workload

N GPUs: every rank runs the same workload on its own stream of pictures (one stream per GPU,
BASELINE configs[4]); no data-path collective; value = N x 120 frames / max-over-ranks time.

`--impl reference` times only the CPU reference (all host cores) on the same metric.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
N_PICS = 121                      # POC 0 (intra) + 15 mini-GOPs of 8
MINIGOP = [(8, 0, 0, 8), (4, 1, -4, 4), (2, 2, -2, 2), (6, 2, -2, 2), (1, 3, -1, 1), (3, 3, -1, 1), (5, 3, -1, 1), (7, 3, -1, 1)]
SB_PER_FRAME = ((W + 63) // 64) * ((H + 63) // 64)


def hme_byte_sads_per_sb_list(prm):
    """Nominal byte-SADs of the HME levels of one full-height SB and one list (search sizes before picture-edge clipping):
    level 0 = 16x16 block of the sixteenth picture, level 1 = 32x32 of the quarter picture, level 2 = 64x64, even rows only."""
    def rnd(w):
        return 8 if w < 8 else (w + (w & 7)) if (w & 7) else w
    total = 0
    for rw in range(prm.num_hme_regions_w):
        for rh in range(prm.num_hme_regions_h):
            if prm.enable_hme_level0:
                total += (prm.hme_l0_w[rw] * prm.hme_l0_mult_x // 100) * (prm.hme_l0_h[rh] * prm.hme_l0_mult_y // 100) * 16 * 8
            if prm.enable_hme_level1:
                total += rnd(prm.hme_l1_w[rw]) * prm.hme_l1_h[rh] * 32 * 16
            if prm.enable_hme_level2:
                total += rnd(prm.hme_l2_w[rw]) * prm.hme_l2_h[rh] * 64 * 32
    return total


USE_16BIT = False                 # me4k: pictures arrive as 16-bit samples (10-bit), split on the device
SHARD_SB_ROWS = False             # me4k: N ranks share every picture by SB rows (strong scaling)
NUM_PUS = 85                      # me209: every picture searched with all 209 PUs (PIC_ALL_DEPTH_MODE, enc-mode 0)
WITH_TXQ = False                  # streams: every frame also gets its luma residual plane transformed + quantised
SEARCH_AREA = None                # --search-area N: full-pel search area N x N instead of the preset's 64 x 64 (the reference searches MIN(N, 127))


def apply_search_area(prm):
    if SEARCH_AREA:
        prm.search_area_w = prm.search_area_h = SEARCH_AREA
    return prm


def fullpel_byte_sads_per_sb_list():
    """algorithmic work of the full-pel kernel per SB and list: positions x (64 x 32 sub-sampled pixels)"""
    n = min(SEARCH_AREA or 64, 127)
    return n * n * 64 * 32


def select_workload(name):
    """me: BASELINE configs[1] (1080p 8-bit, one stream per GPU). me4k: configs[2] (3840x2160 10-bit, SB rows sharded)."""
    global W, H, N_PICS, SB_PER_FRAME, USE_16BIT, SHARD_SB_ROWS, NUM_PUS, WITH_TXQ
    if name == "streams":
        WITH_TXQ = True
    if name == "me209":
        NUM_PUS = 209
    if name == "me4k":
        W, H, N_PICS, USE_16BIT, SHARD_SB_ROWS = 3840, 2160, 33, True, True
        SB_PER_FRAME = ((W + 63) // 64) * ((H + 63) // 64)


def ois_specs(pkg, jobs, rows, first=0, count=None):
    """svtme_ois specs of the bench's pictures: (slot, parameters, ME SADs of the 85 square PUs) per job. `rows`: the host
    me_results images [job][sb][pu] (PU_RESULT_DTYPE). All pictures are B pictures of a 4-layer mini-GOP at enc-mode 3."""
    res = 3 if W >= 3840 else 2 if W >= 1920 else 1 if W >= 1280 else 0
    specs = []
    for j in range(first, len(jobs) if count is None else min(len(jobs), first + count)):
        poc, tl = jobs[j][0], jobs[j][1]
        me = np.ascontiguousarray(rows[j]["dist0"][:, :85]).astype(np.uint32)
        specs.append((poc, pkg.SvtMeOisParams(2, tl, 1 if tl < 3 else 0, 0, res, 0, 3), me))
    return specs


def ois_cpu_ms(frames, jobs, rows):
    """The CPU side of the `ois` key: the reference's own OpenLoopIntraSearchLcu over one layer-0 and one layer-2 picture of the
    sequence (oracle/_ref when it was compiled, else the oracle port), one core, ms per picture."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import me_common as M
    kind = "reference" if cpu_reference_available() else "port"
    if kind == "port":
        M.build_oracle(ref=False)
    lib, pre = (M.load_ref(), "ref") if kind == "reference" else (M.load_oracle(), "meo")
    res = 3 if W >= 3840 else 2 if W >= 1920 else 1 if W >= 1280 else 0
    ms = {}
    for name, want_tl in (("layer0", 0), ("layer2", 2)):
        j = next(i for i, job in enumerate(jobs) if job[1] == want_tl)
        poc, tl = jobs[j][0], jobs[j][1]
        pyr = M.build_pyramid(M.load_oracle(), "meo", np.ascontiguousarray(frames[poc]))
        me = np.ascontiguousarray(rows[j]["dist0"][:, :85]).astype(np.uint32)
        t0 = time.perf_counter()
        M.ois_picture(lib, pre, M.MeoOisParams(W, H, 2, tl, 1 if tl < 3 else 0, 0, res, 0, 3), pyr.c.full, me)
        ms[name] = (time.perf_counter() - t0) * 1e3
    return {"kind": kind, "cores": 1, "ms_per_picture": ms}


def measure_ois(pkg, ctx, jobs, out, frames=None):
    rows = out.numpy().view(pkg.PU_RESULT_DTYPE)[..., 0]
    batches = [ois_specs(pkg, jobs, rows, g, 8) for g in range(0, len(jobs), 8)]
    import torch
    n_sb = rows.shape[1]
    pin = (torch.empty((8, n_sb, 85, 18), dtype=torch.int32).pin_memory().numpy().view(np.uint32),      # results into pinned memory, as e2e does
           torch.empty((8, n_sb, 85), dtype=torch.uint8).pin_memory().numpy())
    ctx.ois(batches[0], out=pin)                                 # warm-up (grows the device buffers once)
    ctx.set_profiling(True)
    ctx.kernel_times(reset=True)
    hist = np.zeros(19, np.int64)
    dt = 0.0
    for b in batches:
        t0 = time.perf_counter()
        _, total = ctx.ois(b, out=pin)
        dt += time.perf_counter() - t0
        hist += np.bincount(total.ravel(), minlength=19)
    ms, n = ctx.kernel_times(reset=True)["ois"]
    ctx.set_profiling(False)
    cpu = None
    if frames is not None:
        try:
            cpu = ois_cpu_ms(frames, jobs, rows)
        except Exception as e:                                   # noqa: BLE001
            cpu = {"error": repr(e)}
    return {"pictures": len(jobs), "launches": int(n), "kernel_us_per_picture": ms / len(jobs) * 1e3,
            "call_us_per_picture": dt / len(jobs) * 1e6, "cpu": cpu,
            "total_intra_luma_mode_histogram": {str(k): int(v) for k, v in enumerate(hist) if v},
            "note": "kernel: CUDA events around the launches (one per mini-GOP of 8 pictures: one layer-0 picture with all 35 modes, "
                    "seven with DC + 3..9 directions by OIS point); call: host wall clock of the synchronous calls incl. ME SADs up and "
                    "3.1 MB of candidate words per picture back into pinned memory"}


def gop_jobs(n_pics=None):
    """[(poc, temporal_layer, ref0_poc, ref1_poc)] for pictures 1..120 in decode order."""
    jobs = []
    for g in range(((n_pics or N_PICS) - 1) // 8):
        base = g * 8
        for off, tl, d0, d1 in MINIGOP:
            poc = base + off
            if tl == 0:
                jobs.append((poc, 0, poc - 8, poc - 8))          # base layer: both lists point at the previous base picture
            else:
                jobs.append((poc, tl, poc + d0, poc + d1))
    return jobs


def synth_sequence(n, seed, w=None, h=None):
    """n luma planes of a panning, noisy synthetic scene (uint8, HxW)."""
    W, H = (w or globals()["W"]), (h or globals()["H"])
    rng = np.random.default_rng(seed)
    big = rng.integers(0, 256, size=((H + 512) // 8, (W + 768) // 8), dtype=np.uint8).astype(np.float32)
    big = np.kron(big, np.ones((8, 8), dtype=np.float32))
    k = np.array([1, 4, 6, 4, 1], dtype=np.float32) / 16.0
    for _ in range(2):
        big = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, big)
        big = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 0, big)
    big = big + rng.normal(0, 10.0, size=big.shape).astype(np.float32)
    out = np.empty((n, H, W), dtype=np.uint8)
    for t in range(n):
        ox, oy = 64 + (3 * t) % 512, 64 + (2 * t) % 256
        f = big[oy:oy + H, ox:ox + W] + rng.normal(0, 2.0, size=(H, W)).astype(np.float32)
        out[t] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return out


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks and throttle reasons while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) == 6:
                    self.rows.append(parts)
            except Exception:
                pass
            self._stop_evt.wait(0.05)                    # nvidia-smi itself takes ~30 ms: ~12 samples per second

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------------
# CPU reference arm: the reference's own ME (oracle/_ref) or, where it is absent, the oracle port.
_CPU = {}


def _cpu_init(frames, kind):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import me_common as M
    _CPU["M"] = M
    _CPU["kind"] = kind
    _CPU["lib"] = M.load_ref() if kind == "reference" else M.load_oracle()
    _CPU["prefix"] = "ref" if kind == "reference" else "meo"
    _CPU["frames"] = frames
    _CPU["pyr"] = {}


def _cpu_task(job):
    """One picture of the hot path on one core: pyramids of the pictures not seen yet by this worker
    (PictureAnalysis decimation + padding) + MotionEstimateLcu for every SB."""
    M, lib, pre = _CPU["M"], _CPU["lib"], _CPU["prefix"]
    poc, tl, r0, r1 = job
    for i in (poc, r0, r1):
        if i not in _CPU["pyr"]:
            _CPU["pyr"][i] = M.build_pyramid(lib, pre, _CPU["frames"][i])
    prm = M.default_params(M.load_oracle(), "meo", W, H, 2, tl, 3, 1 if tl < 3 else 0)
    prm.same_ref_poc = 1 if r0 == r1 else 0
    prm.num_pus = NUM_PUS
    apply_search_area(prm)
    res, _ = M.motion_estimate(lib, pre, prm, _CPU["pyr"][poc], _CPU["pyr"][r0], _CPU["pyr"][r1], want_debug=False)
    return zlib.crc32(res.tobytes())                      # the whole me_results image of the picture (40-byte rows)


def cpu_reference_available():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsvtref.so"))


class CpuArm:
    """Pool of worker processes (one per host core) running the CPU ME on a bounded sample."""

    def __init__(self, frames17):
        import multiprocessing as mp
        self.kind = "reference" if cpu_reference_available() else "port"
        if self.kind == "port":
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import me_common as M
            M.build_oracle(ref=False)
        self.cores = len(os.sched_getaffinity(0))
        self.jobs16 = [j for j in gop_jobs() if j[0] <= 16]
        ctx = mp.get_context("fork")                       # forked before any CUDA context exists in this process
        self.pool = ctx.Pool(self.cores, initializer=_cpu_init, initargs=(frames17, self.kind))

    def run(self, n_tasks):
        tasks = [self.jobs16[i % 16] for i in range(n_tasks)]
        t0 = time.perf_counter()
        crcs = self.pool.map(_cpu_task, tasks, chunksize=1)
        self.crc = {tasks[i][0]: crcs[i] for i in range(min(16, n_tasks))}     # poc -> CRC of the reference's me_results image
        return time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()


def run_reference_impl(args, rank):
    """`--impl reference`: the reference's own CPU implementation of the path, all host threads."""
    if rank != 0:
        return
    frames = synth_sequence(17, seed=1)
    arm = CpuArm(frames)
    if arm.kind == "reference":                            # the workers run it; mapping it here too lets the driver's record of this
        sys.path.insert(0, os.path.join(ROOT, "tests"))    # process's libraries show which CPU implementation the arm timed
        import me_common as M_
        M_.load_ref()
    n_tasks = max(32, 8 * arm.cores)
    for _ in range(args.warmup):
        arm.run(arm.cores)                                 # warm-up also fills each worker's pyramid cache
    times = [arm.run(n_tasks) for _ in range(args.steps)]
    arm.close()
    total = sum(times)
    fps = n_tasks * args.steps / total
    sample = f"{n_tasks} pictures per step drawn from the first two mini-GOPs of the 1080p sequence, one picture per task over {arm.cores} worker processes"
    if WITH_TXQ:                                           # streams: the same cores then transform + quantise one residual plane per frame
        tarm = TxqCpuArm()
        q10c = txq_quant_params()
        sizes = [(w_, h_, 0) for (w_, h_) in TXQ_SIZES]
        tarm.run(sizes[0], q10c, tarm.cores)
        t0, pix = time.perf_counter(), 0
        for _ in range(args.steps):
            for c in sizes:
                tarm.run(c, q10c, tarm.cores)
                pix += tarm.cores * TXQ_CPU_REPS * 128 * TXQ_W
        dt = time.perf_counter() - t0
        tarm.close()
        txq_fps = pix / dt / (TXQ_W * TXQ_H)
        fps = 1.0 / (1.0 / fps + 1.0 / txq_fps)
        total = n_tasks * args.steps / fps
        sample += f"; plus {len(sizes)} transform sizes x {tarm.cores} x {TXQ_CPU_REPS} strips of 1920x128 residual per step ({txq_fps:.0f} planes/s); frame time = ME time + transform time"
    line = {
        "impl": "reference", "metric": "2160p_me_frames_per_sec" if SHARD_SB_ROWS else "1080p_stream_frames_per_sec" if WITH_TXQ else "1080p_me_frames_per_sec", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "strong" if SHARD_SB_ROWS else "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": arm.cores, "kind": arm.kind, "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "sb_per_sec": fps * SB_PER_FRAME, "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def search_area_note():
    if not SEARCH_AREA:
        return "64x64 (preset tables)"
    n = min(SEARCH_AREA, 127)
    return f"{n}x{n} (--search-area {SEARCH_AREA}; the reference searches MIN(area, 127)): {n * n / 4096:.2f}x the algorithmic full-pel work of 64x64"


def workload_config(n_gpus):
    if SHARD_SB_ROWS:
        return {"workload": "3840x2160 10-bit luma (16-bit samples, split 8+2 on upload; ME on the 8-bit plane), 33 pictures (1 intra + 32 B), "
                            "hierarchical mini-GOP 8 / 4 temporal layers, full open-loop ME per picture with the 4K search tables",
                "frames_per_step": N_PICS - 1, "sb_per_frame": SB_PER_FRAME, "lists_per_frame": 2, "pus_per_sb": 85,
                "search_area": search_area_note(),
                "parallelism": f"every picture on every GPU, SB rows of each picture sharded x{n_gpus}, no collective",
                "l2_policy": "inputs larger than L2 (33 planes = 274 MB + 8-bit pyramids per step)"}
    return {"workload": "1920x1080 8-bit luma, 121 pictures (1 intra + 120 B), hierarchical mini-GOP 8 / 4 temporal layers, "
                        f"full open-loop ME per picture: pyramid + HME L0/L1/L2 + 64x64 full-pel ({NUM_PUS} PUs) + half/quarter-pel + bi-pred"
                        + (" [209-PU mode: every picture as an enc-mode 0 PIC_ALL_DEPTH_MODE picture]" if NUM_PUS == 209 else "")
                        + (" + forward transform and quantisation of one 1920x1088 luma residual plane per frame (the 19 DCT_DCT sizes in turn, "
                           "8 planes per launch) on a second stream of the same GPU [BASELINE configs[4]: one such stream per GPU]" if WITH_TXQ else ""),
            "frames_per_step_per_gpu": N_PICS - 1, "sb_per_frame": SB_PER_FRAME, "lists_per_frame": 2, "pus_per_sb": NUM_PUS,
            "search_area": search_area_note(), "streams": n_gpus, "parallelism": f"one picture stream per GPU x{n_gpus}, no collective",
            "l2_policy": "inputs larger than L2 (121 luma planes = 251 MB + 306 MB of padded pyramids per step)"}


# ------------------------------------------------------------------------------------------------
# Sharded modes (north_star: "pictures shard by picture, and by SB row within 4K frames, across the GPUs of one box;
# references replicated peer-to-peer over NVLink; no NCCL collective"). Measured OUTSIDE the headline's timed region and
# reported as extra keys of the default line:
#   strong_1080p  ONE 120-picture 1080p stream dealt to the N ranks (sharding.picture_shard_contiguous); every picture
#                 is uploaded from the host by its owner only and imported by the ranks that reference it straight from the
#                 owner's HBM (svtme_slot_export / svtme_picture_upload_ipc); results return to host memory
#   strong_4k     3840x2160 10-bit pictures, every picture on every GPU, its 2040 SBs sharded in raster order
#                 (SvtMeJob.range_in_sbs, sharding.sb_shard)
#   peer_parity   the sharded 1080p run's me_results of ALL 120 pictures against the single-GPU run of the same stream
#                 (CRC per picture) + the pyramid of one imported picture against a local upload
# Both carry an in-run single-GPU baseline (rank 0 alone runs the whole job while the other ranks wait), so
# efficiency = T(1 GPU) / (N x T(N GPUs)) comes from one process set on one box.
def _timed_steps(torch, ctx, fn, steps, sync_all):
    st = torch.cuda.ExternalStream(ctx.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record(st)
    for _ in range(steps):
        fn()
    ctx.join()
    e1.record(st)
    sync_all()
    return e0.elapsed_time(e1)


def _max_over_ranks(torch, dist, ms):
    if dist is None:
        return ms
    t = torch.tensor([ms], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def measure_strong_1080p(pkg, torch, dist, host_pg, rank, world, local_rank, host, steps, warmup):
    """host: pinned uint8 tensor [N_PICS, H, W] holding the SAME sequence on every rank."""
    shard = importlib.import_module("svt-av1-1_b200.sharding")
    jobs = gop_jobs()
    ctx = pkg.MeContext(W, H, 2 * N_PICS, device=local_rank)
    host_ptrs = [host[i].data_ptr() for i in range(N_PICS)]
    count = [0]

    def sync_all():
        ctx.synchronize()
        torch.cuda.synchronize()

    def plan(my_jobs):
        specs = []
        for poc, tl, r0, r1 in my_jobs:
            p = pkg.default_params(W, H, 2, tl, 3, 1 if tl < 3 else 0)
            p.same_ref_poc = 1 if r0 == r1 else 0
            specs.append((poc, r0, r1, p))
        res = torch.zeros((len(my_jobs), ctx.n_sb, 85, 40), dtype=torch.uint8).pin_memory()
        ptrs = [res[i].data_ptr() for i in range(len(my_jobs))]
        batches = [[ctx.make_jobs([(c + o, a + o, b + o, p) for (c, a, b, p) in specs[g:g + 8]], out_ptrs=ptrs[g:g + 8])[0]
                    for g in range(0, len(specs), 8)] for o in (0, N_PICS)]
        return res, batches

    def crcs(my_jobs, res):
        ctx.synchronize()
        return {j[0]: zlib.crc32(res[i].numpy().tobytes()) for i, j in enumerate(my_jobs)}

    # ---- rank 0 alone: the whole stream on one GPU
    alone_ms, alone_crc = None, None
    if rank == 0:
        res_all, batches_all = plan(jobs)

        def step_alone():
            base = (count[0] & 1) * N_PICS
            count[0] += 1
            for g in range(0, N_PICS - 1, 8):
                pics = list(range(0 if g == 0 else g + 1, g + 9))
                ctx.upload_ptr_batch([base + i for i in pics], [host_ptrs[i] for i in pics], W)
            for jb in batches_all[(count[0] - 1) & 1]:
                ctx.motion_estimate_async(jb)
        for _ in range(max(3, warmup)):                   # >= 3: every entry of the context's batch ring allocates on first use
            step_alone()
        alone_ms = _timed_steps(torch, ctx, step_alone, steps, sync_all) / steps
        alone_crc = crcs(jobs, res_all)
    if world == 1:
        ctx.close()
        return {"n_gpus": 1, "frames_per_sec": (N_PICS - 1) / (alone_ms / 1e3), "ms_per_step": alone_ms, "efficiency": 1.0}, "n/a (one GPU)"
    dist.barrier(group=host_pg)

    # ---- all ranks: the stream dealt by picture, references imported from their owners
    mine, need, owner = shard.picture_shard_contiguous(jobs, world, rank)
    own = [p for p in range(N_PICS) if owner.get(p) == rank]
    imports = [p for p in need if owner[p] != rank]
    tokens = [None] * world
    dist.all_gather_object(tokens, [ctx.export_slot(i) for i in range(2 * N_PICS)], group=host_pg)
    res_mine, batches_mine = plan(mine)
    count[0] = 0

    def fetch_pictures(which):
        """pictures of one sequence into slot set `which`: my own from the host, the others' from their owners' HBM"""
        base = which * N_PICS
        ctx.synchronize_uploads()                       # my imports of the previous sequence have landed ...
        for i0 in range(0, len(own), 64):
            ctx.upload_ptr_batch([base + i for i in own[i0:i0 + 64]], [host_ptrs[i] for i in own[i0:i0 + 64]], W)
        dist.barrier(group=host_pg)                     # ... on every rank, and every owner has ISSUED this sequence's uploads
        for i in imports:                               # ordered behind the owner's upload on the device (inter-process event)
            ctx.upload_ipc(base + i, tokens[owner[i]][base + i])

    def step_sharded():
        # software pipeline, as a streaming encoder's look-ahead would run it: the NEXT sequence's pictures are fetched (host
        # copies, barrier, NVLink imports) while this sequence is searched; every timed step holds one fetch and one search
        which = count[0] & 1
        count[0] += 1
        for jb in batches_mine[which]:                  # searches first: the GPU has work queued while the host sits in the barrier
            ctx.motion_estimate_async(jb)
        fetch_pictures(which ^ 1)

    def step_search_only():                             # diagnostic: the same batches with the pictures left in their slots
        which = count[0] & 1
        count[0] += 1
        for jb in batches_mine[which]:
            ctx.motion_estimate_async(jb)
    fetch_pictures(0)

    def sync_ranks():
        sync_all()
        dist.barrier(group=host_pg)
    for _ in range(max(3, warmup)):
        step_sharded()
    ms = _max_over_ranks(torch, dist, _timed_steps(torch, ctx, step_sharded, steps, sync_ranks) / steps)
    ms_search = _max_over_ranks(torch, dist, _timed_steps(torch, ctx, step_search_only, steps, sync_ranks) / steps)
    # ---- parity: every rank's rows against rank 0's single-GPU run; one imported picture's pyramid against a local upload
    my_crc = crcs(mine, res_mine)
    pyr_ok = True
    if imports:
        i = imports[0]
        last = ((count[0] - 1) & 1) * N_PICS           # slot set of the last step
        other = N_PICS - last
        spare = [p for p in range(N_PICS) if p not in need][0]      # a slot no job of this rank reads
        ctx.upload_ptr(other + spare, host_ptrs[i], W)
        ctx.synchronize()
        pyr_ok = all(np.array_equal(ctx.download_plane(last + i, lvl), ctx.download_plane(other + spare, lvl)) for lvl in range(3))
    gathered = [None] * world
    dist.all_gather_object(gathered, (my_crc, pyr_ok, len(own), len(imports)), group=host_pg)
    dist.barrier(group=host_pg)
    ctx.close()
    if rank != 0:
        return None, None
    bad = [poc for g in gathered for poc, c in g[0].items() if alone_crc[poc] != c]
    n_checked = sum(len(g[0]) for g in gathered)
    parity = ("ok: %d pictures' me_results identical to the single-GPU run, imported pyramids identical" % n_checked
              if not bad and all(g[1] for g in gathered) and n_checked == len(jobs) else "MISMATCH: pictures %s, pyramids %s" % (bad[:8], [g[1] for g in gathered]))
    return {"n_gpus": world, "frames_per_sec": (N_PICS - 1) / (ms / 1e3), "ms_per_step": ms,
            "single_gpu_frames_per_sec": (N_PICS - 1) / (alone_ms / 1e3), "single_gpu_ms_per_step": alone_ms,
            "efficiency": alone_ms / (world * ms),
            "search_only_ms_per_step": ms_search, "search_only_efficiency": alone_ms / (world * ms_search),
            "search_only_note": "the same per-rank batches with no fetch (pictures left in their slots, results still copied out): "
                                "what the smaller per-rank batches cost on the GPU; the rest of the gap is the fetch (host copies, barrier, imports)",
            "pictures_per_rank": [len(g[0]) for g in gathered], "uploaded_per_rank": [g[2] for g in gathered],
            "imported_over_nvlink_per_rank": [g[3] for g in gathered],
            "h2d_bytes_per_step_per_rank": max(g[2] for g in gathered) * W * H,
            "what": "one 120-picture 1080p stream, pictures dealt to the ranks in contiguous runs of the decode order; host buffers in, "
                    "me_results to host memory out, inside the timed region; one host barrier per step, no data-path collective"}, parity


def measure_strong_4k(pkg, torch, dist, host_pg, rank, world, local_rank, steps, warmup):
    shard = importlib.import_module("svt-av1-1_b200.sharding")
    w4, h4, n_pics = 3840, 2160, 33
    frames = synth_sequence(n_pics, seed=1, w=w4, h=h4)
    lsb = np.random.default_rng(5).integers(0, 4, size=frames.shape, dtype=np.uint16)
    host16 = torch.from_numpy((frames.astype(np.uint16) << 2) | lsb).pin_memory()
    dev8 = torch.from_numpy(frames).cuda()
    del frames, lsb
    jobs = gop_jobs(n_pics)
    ctx = pkg.MeContext(w4, h4, 2 * n_pics, device=local_rank)
    dev_ptrs = [dev8[i].data_ptr() for i in range(n_pics)]
    specs = []
    for poc, tl, r0, r1 in jobs:
        p = pkg.default_params(w4, h4, 2, tl, 3, 1 if tl < 3 else 0)
        p.same_ref_poc = 1 if r0 == r1 else 0
        specs.append((poc, r0, r1, p))
    count = [0]

    def sync_all():
        ctx.synchronize()
        torch.cuda.synchronize()

    def sync_ranks():
        sync_all()
        if dist is not None:
            dist.barrier(group=host_pg)

    def plan(sb_range):
        n_sb = ctx.n_sb if sb_range is None else sb_range[1] - sb_range[0]
        res = torch.zeros((len(jobs), n_sb, 85, 40), dtype=torch.uint8).pin_memory()
        ptrs = [res[i].data_ptr() for i in range(len(jobs))]
        resident = ctx.make_jobs(specs, out_ptrs=[None] * len(jobs), sb_range=sb_range)[0]
        e2e = [[ctx.make_jobs([(c + o, a + o, b + o, p) for (c, a, b, p) in specs[g:g + 8]], out_ptrs=ptrs[g:g + 8], sb_range=sb_range)[0]
                for g in range(0, len(specs), 8)] for o in (0, n_pics)]
        return res, resident, e2e

    def make_steps(resident, e2e):
        def step_resident():
            ctx.upload_device_batch(list(range(n_pics)), dev_ptrs, w4)
            ctx.motion_estimate_async(resident)

        def step_e2e():
            which = count[0] & 1
            count[0] += 1
            for i in range(n_pics):                    # every rank needs every picture: each uploads all of them over its own host link
                rc = ctx.lib.svtme_picture_upload_16bit(ctx._h, which * n_pics + i, host16[i].data_ptr(), w4, None)
                assert rc == 0
            for jb in e2e[which]:
                ctx.motion_estimate_async(jb)

        def fetch_replicated(which):
            """each picture crosses the host link ONCE (owner = picture index mod N); the other ranks read it from the owner's
            HBM over NVLink while building their own pyramid (svtme_picture_upload_ipc)"""
            base = which * n_pics
            ctx.synchronize_uploads()                   # my imports of the previous sequence have landed ...
            for i in range(rank, n_pics, world):
                rc = ctx.lib.svtme_picture_upload_16bit(ctx._h, base + i, host16[i].data_ptr(), w4, None)
                assert rc == 0
            dist.barrier(group=host_pg)                 # ... on every rank, and every owner has ISSUED this sequence's uploads
            for i in range(n_pics):
                if i % world != rank:
                    ctx.upload_ipc(base + i, tokens[i % world][base + i])

        def step_e2e_replicated():
            # software pipeline: fetch the next sequence's pictures while this one is searched (one fetch + one search per step)
            which = count[0] & 1
            count[0] += 1
            for jb in e2e[which]:
                ctx.motion_estimate_async(jb)
            fetch_replicated(which ^ 1)
        step_e2e_replicated.prime = lambda: fetch_replicated(0)
        return step_resident, step_e2e, step_e2e_replicated

    alone = None
    if rank == 0:
        _, resident, e2e = plan(None)
        sr, se, _ = make_steps(resident, e2e)
        for _ in range(max(3, warmup)):
            sr(); se()
        alone = (_timed_steps(torch, ctx, sr, steps, sync_all) / steps, _timed_steps(torch, ctx, se, steps, sync_all) / steps)
    out = {"n_gpus": world, "frames_per_step": n_pics - 1, "sb_per_frame": ctx.n_sb,
           "what": "3840x2160 10-bit (16-bit samples split on upload), 33 pictures, every picture on every GPU, the 2040 SBs of each picture "
                   "sharded in raster order (SvtMeJob.range_in_sbs); HME level 2 off as in the enc-mode 3 4K tables"}
    if world == 1:
        ctx.close()
        out.update({"resident": {"frames_per_sec": (n_pics - 1) / (alone[0] / 1e3), "ms_per_step": alone[0], "efficiency": 1.0},
                    "e2e": {"frames_per_sec": (n_pics - 1) / (alone[1] / 1e3), "ms_per_step": alone[1], "efficiency": 1.0}})
        return out
    dist.barrier(group=host_pg)
    first, cnt = shard.sb_shard(ctx.n_sb, world, rank)
    res_mine, resident, e2e = plan((first, first + cnt))
    tokens = [None] * world
    dist.all_gather_object(tokens, [ctx.export_slot(i) for i in range(2 * n_pics)], group=host_pg)
    sr, se, sx = make_steps(resident, e2e)
    count[0] = 0
    for _ in range(max(3, warmup)):
        sr(); se()
    ms_r = _max_over_ranks(torch, dist, _timed_steps(torch, ctx, sr, steps, sync_ranks) / steps)
    ms_e = _max_over_ranks(torch, dist, _timed_steps(torch, ctx, se, steps, sync_ranks) / steps)
    ctx.synchronize()
    crc_own = zlib.crc32(res_mine.numpy().tobytes())    # rows from independently uploaded pictures
    count[0] = 0
    res_mine.zero_()
    sx.prime()
    for _ in range(3):
        sx()
    ms_x = _max_over_ranks(torch, dist, _timed_steps(torch, ctx, sx, steps, sync_ranks) / steps)
    ctx.synchronize()
    same = [None] * world
    dist.all_gather_object(same, zlib.crc32(res_mine.numpy().tobytes()) == crc_own, group=host_pg)
    dist.barrier(group=host_pg)
    ctx.close()
    if rank != 0:
        return None
    out["sbs_per_rank"] = cnt
    out["resident"] = {"frames_per_sec": (n_pics - 1) / (ms_r / 1e3), "ms_per_step": ms_r, "single_gpu_ms_per_step": alone[0],
                       "efficiency": alone[0] / (world * ms_r)}
    out["e2e"] = {"frames_per_sec": (n_pics - 1) / (ms_e / 1e3), "ms_per_step": ms_e, "single_gpu_ms_per_step": alone[1],
                  "efficiency": alone[1] / (world * ms_e), "h2d_bytes_per_step_per_rank": n_pics * w4 * h4 * 2,
                  "note": "every rank uploads every picture over its own host link: the upload does not shrink with N"}
    out["e2e_replicated"] = {"frames_per_sec": (n_pics - 1) / (ms_x / 1e3), "ms_per_step": ms_x, "single_gpu_ms_per_step": alone[1],
                             "efficiency": alone[1] / (world * ms_x), "h2d_bytes_per_step_per_rank": len(range(0, n_pics, world)) * w4 * h4 * 2,
                             "rows_identical_to_independent_uploads": all(same),
                             "note": "each picture uploaded by one rank (index mod N) and read by the others from that rank's HBM over NVLink "
                                     "(svtme_picture_upload_ipc); one host barrier per step"}
    return out


# ------------------------------------------------------------------------------------------------
# GPU arm
def int_pipe_peak(sm_mhz):
    """Peak of the binding unit for the full-pel kernel: VABSDIFF4.U8.ACC issues at 64 lanes/clk/SM on
    sm_100a (measured: profiles/microbench/int_pipe_bench_r01.jsonl, 122.5 lane-ops/ns/SM at 1.965 GHz
    = 62.3 lanes/clk), 4 byte-SADs per lane-op, 148 SMs. Returned in tera byte-SADs per second at the
    given clock (the clock observed under load, else the maximum)."""
    return 148 * 64 * 4 * sm_mhz * 1e6 / 1e12


def run_gpu_impl(args, rank, world, local_rank):
    import torch
    pkg = importlib.import_module("svt-av1-1_b200")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # ---- CPU baseline first (rank 0, N=1 only), while this process has no CUDA context yet
    cpu, cpu_crc = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        frames17 = synth_sequence(17, seed=1)
        arm = CpuArm(frames17)
        n_tasks = max(64, 24 * arm.cores)                  # ~6 s wall of the reference's ME on every host core
        arm.run(arm.cores)
        t = arm.run(n_tasks)
        arm.close()
        cpu_crc = dict(arm.crc)                            # CRC of the CPU implementation's me_results image, pictures 1..16
        cpu = {"value": n_tasks / t, "unit": "frames/s", "cores": arm.cores, "kind": arm.kind,
               "sample": f"{n_tasks} pictures of the first two mini-GOPs, {arm.cores} worker processes, {t:.1f} s wall"}
        if WITH_TXQ:                                       # the same cores then transform + quantise: time per frame adds up
            tarm = TxqCpuArm()
            q10c = txq_quant_params()
            sizes = [(w_, h_, 0) for (w_, h_) in TXQ_SIZES]
            tarm.run(sizes[0], q10c, tarm.cores)
            t0 = time.perf_counter()
            pix = 0
            for c in sizes:
                tarm.run(c, q10c, 2 * tarm.cores)
                pix += 2 * tarm.cores * TXQ_CPU_REPS * 128 * TXQ_W
            dt = time.perf_counter() - t0
            tarm.close()
            txq_fps = pix / dt / (TXQ_W * TXQ_H)
            cpu["me_frames_per_sec"], cpu["txq_frames_per_sec"] = cpu["value"], txq_fps
            cpu["value"] = 1.0 / (1.0 / cpu["value"] + 1.0 / txq_fps)
            cpu["sample"] += f"; then {len(sizes)} transform sizes x {2 * tarm.cores} x {TXQ_CPU_REPS} strips of 1920x128 residual, {dt:.1f} s wall"

    torch.cuda.set_device(local_rank)
    frames = synth_sequence(N_PICS, seed=1 if SHARD_SB_ROWS else 1 + rank)   # row sharding: every rank sees the same pictures
    if USE_16BIT:
        lsb = np.random.default_rng(5).integers(0, 4, size=frames.shape, dtype=np.uint16)
        host = torch.from_numpy((frames.astype(np.uint16) << 2) | lsb).pin_memory()   # pinned 16-bit planes (e2e source)
        dev = torch.from_numpy(frames).cuda()                                          # 8-bit planes resident in HBM (value source)
    else:
        host = torch.from_numpy(frames).pin_memory()                   # pinned host planes (e2e source)
        dev = host.cuda(non_blocking=False)                             # HBM-resident planes (value source)
    jobs = gop_jobs()
    # two sets of picture slots: consecutive e2e steps (consecutive sequences of a stream) alternate between them, so
    # the uploads of one sequence overlap the searches of the previous one, as in a streaming encoder with look-ahead
    ctx = pkg.MeContext(W, H, 2 * N_PICS, device=local_rank)
    specs = []
    for poc, tl, r0, r1 in jobs:
        p = pkg.default_params(W, H, 2, tl, 3, 1 if tl < 3 else 0)
        p.same_ref_poc = 1 if r0 == r1 else 0
        p.num_pus = NUM_PUS
        apply_search_area(p)
        specs.append((poc, r0, r1, p))
    sb_rows = None
    if SHARD_SB_ROWS and world > 1:
        shard = importlib.import_module("svt-av1-1_b200.sharding")
        first, count = shard.sb_row_shard((H + 63) // 64, world, rank)
        sb_rows = (first, first + count)
    n_sb = ctx.n_sb if sb_rows is None else (sb_rows[1] - sb_rows[0]) * ((W + 63) // 64)
    out = torch.empty((len(jobs), n_sb, NUM_PUS, 40), dtype=torch.uint8).pin_memory()     # host me_results images
    ptrs = [out[i].data_ptr() for i in range(len(jobs))]
    job_resident, _, _ = ctx.make_jobs(specs, out_ptrs=[None] * len(jobs), sb_rows=sb_rows)   # results stay in HBM
    # e2e: one batch per few mini-GOPs so that uploads, searches and result copies of consecutive batches overlap on the
    # context's three streams
    EB = 8 * int(os.environ.get("SVTME_BENCH_E2E_GOPS", "1"))     # pictures per e2e batch: whole mini-GOPs (measured: 1, 3, 5 mini-GOPs per batch give 39.5, 40.1, 39.2 ms per step)
    gop_batches = [[ctx.make_jobs([(c + o, r0 + o, r1 + o, p) for (c, r0, r1, p) in specs[g:g + EB]], out_ptrs=ptrs[g:g + EB], sb_rows=sb_rows)[0]
                    for g in range(0, len(jobs), EB)] for o in (0, N_PICS)]
    stride = W
    e2e_count = [0]
    txq = None
    if WITH_TXQ:
        # second kernel family on its own stream of the same GPU: 120 residual planes per step, 8 stacked planes per launch
        tq = {"ctx": pkg.TxqContext(local_rank), "quant": pkg.SvtTxqQuant.from_q10(txq_quant_params()),
              "sizes": [(w_, h_, 0) for (w_, h_) in TXQ_SIZES], "n": 0}
        res = txq_residual(11 + rank, TXQ_PLANES)
        tq["host"] = torch.from_numpy(res).pin_memory()
        tq["dev"] = tq["host"].cuda()
        npix = TXQ_PLANES * TXQ_H * TXQ_W
        tq["d_out"] = [torch.empty(npix, dtype=torch.int32, device="cuda") for _ in range(3)]
        tq["d_eob"] = torch.empty(npix // 16, dtype=torch.int16, device="cuda")
        tq["h_lv"] = torch.empty(npix, dtype=torch.int16).pin_memory()
        tq["h_off"] = torch.empty(npix // 16 + 1, dtype=torch.int32).pin_memory()
        tq["h_eob"] = torch.empty(npix // 16, dtype=torch.int16).pin_memory()
        tq["levels"] = 0
        tq["st"] = torch.cuda.ExternalStream(tq["ctx"].stream())
        tq["me_st"] = torch.cuda.ExternalStream(ctx.stream)
        tq["launches_per_step"] = (N_PICS - 1) // TXQ_PLANES
        txq = tq

    def txq_mark():
        a = torch.cuda.Event()
        a.record(txq["me_st"])
        return a

    def txq_step_resident(a):
        """15 launches x 8 planes on the transform stream, overlapping this step's searches: they start at the ME stream's
        position `a` (taken before the searches were enqueued) and the ME stream waits for them at the end of the step."""
        txq["st"].wait_event(a)
        for _ in range(txq["launches_per_step"]):
            c = txq["sizes"][txq["n"] % len(txq["sizes"])]
            txq["n"] += 1
            txq["ctx"].fwd_txfm_quant_device(txq["dev"].data_ptr(), TXQ_W, TXQ_W, TXQ_PLANES * TXQ_H, c[0], c[1], c[2], txq["quant"],
                                             txq["d_out"][0].data_ptr(), txq["d_out"][1].data_ptr(), txq["d_out"][2].data_ptr(),
                                             txq["d_eob"].data_ptr(), None)
        b = torch.cuda.Event()
        b.record(txq["st"])
        txq["me_st"].wait_event(b)

    def txq_step_e2e(n_calls=None):
        """The same through host buffers: residual planes H2D; eob, level offsets and the eob leading quantised levels of every
        block D2H (svttxq_fwd_txfm_quant_compact: what the entropy coder reads, dqcoeff is rebuilt by the caller); synchronous
        calls on this thread while the ME batches run asynchronously."""
        for _ in range(txq["launches_per_step"] if n_calls is None else n_calls):
            c = txq["sizes"][txq["n"] % len(txq["sizes"])]
            txq["n"] += 1
            n = C.c_size_t(0)
            rc = txq["ctx"].lib.svttxq_fwd_txfm_quant_compact(txq["ctx"]._h, txq["host"].data_ptr(), TXQ_W, TXQ_W, TXQ_PLANES * TXQ_H, c[0], c[1], c[2],
                                                              C.byref(txq["quant"]), txq["h_eob"].data_ptr(), txq["h_off"].data_ptr(),
                                                              txq["h_lv"].data_ptr(), txq["h_lv"].numel(), C.byref(n))
            assert rc == 0
            txq["levels"] += n.value

    dev_ptrs = [dev[i].data_ptr() for i in range(N_PICS)]
    host_ptrs = [host[i].data_ptr() for i in range(N_PICS)]

    def step_resident():
        mark = txq_mark() if txq else None
        # the pyramids of the whole sequence in one launch (svtme_pictures_upload_device; <= 128 pictures per call)
        for i0 in range(0, N_PICS, 128):
            ctx.upload_device_batch(list(range(i0, min(i0 + 128, N_PICS))), dev_ptrs[i0:i0 + 128], stride)
        ctx.motion_estimate_async(job_resident)
        if txq:
            txq_step_resident(mark)

    # Optional: one rank uploads each picture and the others import it over NVLink (svtme_picture_upload_ipc). Off by default:
    # on this box every GPU has its own host link, and the per-step hand-shakes cost more than the saved H2D bytes
    # (N=2, 4K: 1155 frames/s replicated vs 1368 with independent uploads).
    replicate = SHARD_SB_ROWS and world > 1 and os.environ.get("SVTME_BENCH_REPLICATE") == "1"
    if replicate:
        # one rank uploads each picture from the host; the others import it over NVLink (CUDA IPC + peer copy)
        host_pg = dist.new_group(backend="gloo")            # host-side hand-shakes only
        tokens = [None] * world
        dist.all_gather_object(tokens, [ctx.export_slot(i) for i in range(2 * N_PICS)])

    def step_e2e_replicated(base):
        ctx.synchronize_uploads()                           # host waits for earlier copies only; the searches keep running
        dist.barrier(group=host_pg)                         # ... on every rank: importers are done, owners may rewrite slots
        for i in range(rank, N_PICS, world):
            rc = ctx.lib.svtme_picture_upload_16bit(ctx._h, base + i, host[i].data_ptr(), stride, None)
            assert rc == 0
        dist.barrier(group=host_pg)                         # every owner has issued its uploads
        for i in range(N_PICS):
            if i % world != rank:
                ctx.upload_ipc(base + i, tokens[i % world][base + i])

    def step_e2e():
        which = e2e_count[0] & 1
        e2e_count[0] += 1
        base = which * N_PICS
        if replicate:
            step_e2e_replicated(base)
        else:
            if USE_16BIT:
                for i in range(N_PICS):
                    rc = ctx.lib.svtme_picture_upload_16bit(ctx._h, base + i, host[i].data_ptr(), stride, None)
                    assert rc == 0
            else:
                # pictures arrive a few mini-GOPs at a time (picture 0 with the first group): their copies + one pyramid launch
                for g in range(0, N_PICS - 1, EB):
                    pics = list(range(0 if g == 0 else g + 1, min(g + EB, N_PICS - 1) + 1))
                    ctx.upload_ptr_batch([base + i for i in pics], [host_ptrs[i] for i in pics], stride)
        if txq:
            # the transform calls are synchronous, so they run on their own host thread (ctypes releases the GIL) like the EncDec
            # threads of a pipelined encoder, beside this thread's asynchronous ME submissions; joined at the end of the step
            th = threading.Thread(target=txq_step_e2e)
            th.start()
        for jb in gop_batches[which]:
            ctx.motion_estimate_async(jb)
        if txq:
            th.join()

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def timed(fn, steps):
        """Device time of `steps` calls: CUDA events on the context stream (the stream the kernels run on)."""
        st = torch.cuda.ExternalStream(ctx.stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(st)
        for _ in range(steps):
            fn()
        ctx.join()                                      # compute stream now also waits for uploads / result copies
        e1.record(st)
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ctx.set_profiling(not os.environ.get("SVTME_BENCH_NO_KERNEL_TIMES"))    # per-kernel events keep every batch on one stream
    ctx.kernel_times(reset=True)
    l0 = ctx.kernel_launches
    tl0 = txq["ctx"].kernel_launches() if txq else 0
    ms = timed(step_resident, args.steps)
    launches = ctx.kernel_launches - l0
    txq_resident_launches = (txq["ctx"].kernel_launches() - tl0) if txq else 0
    kt = ctx.kernel_times(reset=True)
    ctx.set_profiling(False)
    clocks = sampler.stop() if sampler else None

    # ---- parity of the bench's own workload (outside the timed regions): the me_results images of the first 16 pictures,
    # from the resident path's slots and again after the e2e steps, against the CPU implementation's (the checker, cpu_baseline)
    def check_rows():
        ctx.synchronize()
        return sum(1 for i, j in enumerate(jobs) if j[0] in cpu_crc and zlib.crc32(out[i].numpy().tobytes()) == cpu_crc[j[0]])
    parity_checked = None
    whole_pictures = sb_rows is None
    if cpu_crc and whole_pictures:
        for jb in gop_batches[0][:(16 + EB - 1) // EB]:
            ctx.motion_estimate_async(jb)                  # slots 0..16 as the resident steps left them
        n_res = check_rows()
        out.zero_()
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    ctx.fetch(gop_batches[0][-1])
    if cpu_crc and whole_pictures:
        n_e2e = check_rows()
        n_want = sum(1 for j in jobs if j[0] in cpu_crc)
        parity_checked = {"pictures": n_want, "resident_identical": n_res, "e2e_identical": n_e2e,
                          "against": cpu["kind"] + " (whole 40-byte me_results rows of every SB and PU, CRC per picture)",
                          "ok": n_res == n_want and n_e2e == n_want}
    checksum = int(out.view(torch.int32).sum(dtype=torch.int64).item()) if rank == 0 else 0
    # ---- side product of the ME process (outside the timed regions): zero-motion SADs of the 120 consecutive picture pairs of
    # the resident pictures (svtme_zz_sad = ComputeDecimatedZzSad), one batched launch + 2 x 510 bytes back per pair
    zz_sad = None
    if rank == 0 and not USE_16BIT:
        pairs = [(i, i - 1) for i in range(1, N_PICS)]
        ctx.zz_sad(pairs)
        t0 = time.perf_counter()
        zz, nm = ctx.zz_sad(pairs)
        dt = time.perf_counter() - t0
        zz_sad = {"pairs": len(pairs), "us_per_picture_pair": dt / len(pairs) * 1e6, "launches": 1,
                  "classes": {str(int(k)): int(v) for k, v in zip(*np.unique(zz, return_counts=True))},
                  "note": "host wall clock of one synchronous call for all pairs: one launch (512 bytes read per SB) and one copy of 2 bytes per SB"}
    # ---- second side product: the open-loop intra search of the same 120 pictures (svtme_ois = OpenLoopIntraSearchLcu), one
    # launch per mini-GOP of 8, steered by the ME SADs the e2e steps left in the host rows. Never fatal for the bench line.
    ois = None
    if rank == 0 and not USE_16BIT and whole_pictures:
        try:
            ois = measure_ois(pkg, ctx, jobs, out, None if args.no_cpu_baseline else frames)
        except Exception as e:                                   # noqa: BLE001 -- reported in the line instead
            ois = {"error": repr(e)}
    ctx.close()
    del dev
    strong_1080p = strong_4k = peer_parity = None
    if args.workload == "me" and not SEARCH_AREA and not args.no_strong:
        host_pg = dist.new_group(backend="gloo") if dist is not None else None      # host-side hand-shakes only
        host1 = host if rank == 0 else torch.from_numpy(synth_sequence(N_PICS, seed=1)).pin_memory()   # ONE stream, the same on every rank
        strong_1080p, peer_parity = measure_strong_1080p(pkg, torch, dist, host_pg, rank, world, local_rank, host1, max(2, args.steps // 2), args.warmup)
        del host1
        strong_4k = measure_strong_4k(pkg, torch, dist, host_pg, rank, world, local_rank, max(2, args.steps // 2), args.warmup)

    if rank == 0:
        frames_total = (1 if SHARD_SB_ROWS else world) * (N_PICS - 1) * args.steps
        fps = frames_total / (ms / 1e3)
        fps_e2e = frames_total / (ms_e2e / 1e3)
        fp_ms, fp_n = kt["fullpel"]
        total_kernel_ms = sum(v[0] for v in kt.values())
        # algorithmic work of one full-pel launch: every SB of every picture of the batch, one list
        sads_per_launch = fullpel_byte_sads_per_sb_list() * n_sb * len(jobs)
        achieved = sads_per_launch / (fp_ms / fp_n / 1e3) / 1e12 if fp_n and fp_ms else None
        sb_cols_, n_sb_pic, full_rows_sb = (W + 63) // 64, ((W + 63) // 64) * ((H + 63) // 64), (H // 64) * ((W + 63) // 64)
        step_sads = (1 if SHARD_SB_ROWS else world) * sum(
            sp[3].num_lists * (fullpel_byte_sads_per_sb_list() * n_sb_pic + hme_byte_sads_per_sb_list(sp[3]) * full_rows_sb) for sp in specs)
        traffic = None                                    # DRAM bytes per launch, scaled from the committed ncu capture
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02", "traffic.json")))
            traffic = tj["fullpel_ext_kernel" if NUM_PUS == 209 else "fullpel_kernel"]["dram_bytes_per_sb_list"] * n_sb * len(jobs)
        except Exception:
            pass
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965
        peak = int_pipe_peak(sm_mhz)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        py_ms, py_n = kt["pyramid"]
        txq_h2d = txq_d2h = txq_launches = 0
        if txq:                                           # per step: 15 launches of 8 planes; e2e moves the residual in and qcoeff/dqcoeff/eob out
            plane_px = TXQ_PLANES * TXQ_H * TXQ_W
            txq_h2d = txq["launches_per_step"] * 2 * plane_px
            e2e_calls = max(1, (max(1, args.warmup // 2) + args.steps) * txq["launches_per_step"])
            txq_d2h = txq["launches_per_step"] * (2 * txq["levels"] // e2e_calls) + sum(
                txq["launches_per_step"] * 6 * (plane_px // (w_ * h_)) for (w_, h_, _) in txq["sizes"]) // len(txq["sizes"])
            txq_launches = txq_resident_launches
        pyr_bytes = W * H + sum(((W >> s) + 2 * p) * ((H >> s) + 2 * p) for s, p in ((0, 68), (1, 32), (2, 16)))
        line = {
            "metric": "2160p_me_frames_per_sec" if SHARD_SB_ROWS else "1080p_stream_frames_per_sec" if WITH_TXQ else "1080p_me_frames_per_sec",
            "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if SHARD_SB_ROWS else "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": workload_config(world),
            "sb_per_sec": fps * SB_PER_FRAME,
            "e2e": {"value": fps_e2e, "unit": "frames/s",
                    "h2d_bytes_per_step": (len(range(rank, N_PICS, world)) if replicate else N_PICS) * W * H * (2 if USE_16BIT else 1) + txq_h2d,
                    "replication": "each picture uploaded by one rank, imported by the others over NVLink (svtme_picture_upload_ipc)" if replicate else None,
                    "d2h_bytes_per_step": len(jobs) * n_sb * NUM_PUS * 40 + txq_d2h, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches) + txq_launches,
            "roofline": {"kernel": "fullpel_ext_kernel" if NUM_PUS == 209 else "fullpel_kernel", "bound": "int", "achieved": achieved, "peak": peak, "unit": "Tbyte-SAD/s",
                         "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                         "traffic_source": "dram__bytes_read+write per SB-list from profiles/r02/traffic.json (ncu --set full) x SB-lists per launch; bytes",
                         "peak_source": f"148 SM x 64 lanes/clk x 4 bytes x {sm_mhz} MHz (VABSDIFF4 issue rate, profiles/microbench)",
                         "launch_ms": fp_ms / fp_n if fp_n else None, "share_of_kernel_time": fp_ms / total_kernel_ms if total_kernel_ms else None},
            # one launch builds the pyramids of the whole sequence: bytes of all N_PICS pictures over the launches' time
            "roofline_hbm": {"kernel": "pyramid_flat_kernel", "bound": "hbm", "pictures_per_launch": N_PICS * args.steps / py_n if py_n else None,
                             "achieved": (pyr_bytes * N_PICS * args.steps / (py_ms / 1e3) / 1e9) if py_n and py_ms else None,
                             "peak": hbm_peak, "unit": "GB/s", "frac": (pyr_bytes * N_PICS * args.steps / (py_ms / 1e3) / 1e9 / hbm_peak) if py_n and py_ms else None,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s"},
            # the whole step against the same peak: HME + full-pel byte-SADs of every (picture, SB, list) over the step time
            # (the sub-pel / bi-pred arithmetic is not counted, so this under-states the work done)
            "roofline_step": {"bound": "int", "unit": "Tbyte-SAD/s", "peak": peak * world,
                              "achieved": step_sads / (ms / args.steps / 1e3) / 1e12,
                              "frac": step_sads / (ms / args.steps / 1e3) / 1e12 / (peak * world),
                              "byte_sads_per_step": step_sads},
            "kernel_ms_per_step": {k: v[0] / args.steps for k, v in kt.items()},
            "kernel_launches_per_step": {k: v[1] / args.steps for k, v in kt.items()},
            "clocks": clocks, "result_checksum": checksum, "parity_checked": parity_checked,
            "strong_1080p": strong_1080p, "strong_4k": strong_4k, "peer_parity": peer_parity,
            "zz_sad": zz_sad, "ois": ois,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# Secondary kernel family: forward transform + quantisation sweep (BASELINE.json configs[3]); `--workload txq`
TXQ_SIZES = [(4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (4, 8), (8, 4), (8, 16), (16, 8), (16, 32), (32, 16), (32, 64),
             (64, 32), (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
TXQ_PLANES = 8                                     # 1080p residual planes stacked per launch (outputs 200 MB > L2)
TXQ_W, TXQ_H = 1920, 1088
TXQ_CPU_REPS = 32                                  # repetitions of one 1920x128 strip per CPU task


def txq_cases():
    """(txw, txh, tx_type): DCT_DCT for every size, ADST_ADST and IDTX where the reference has them."""
    out = []
    for (w, h) in TXQ_SIZES:
        out.append((w, h, 0))
        if max(w, h) <= 16:
            out.append((w, h, 3))
        if max(w, h) <= 32:
            out.append((w, h, 9))
    return out


def txq_residual(seed, planes):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:TXQ_H, 0:TXQ_W]
    base = 20 * np.sin(x / 7.3) * np.cos(y / 11.1)
    out = np.empty((planes * TXQ_H, TXQ_W), np.int16)
    for i in range(planes):
        out[i * TXQ_H:(i + 1) * TXQ_H] = np.clip(np.rint(base + rng.normal(0, 8, size=base.shape)), -255, 255).astype(np.int16)
    return out


def _txq_cpu_init(kind):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import txq_common as T
    _CPU["T"] = T
    _CPU["kind"] = kind
    _CPU["lib"] = T.reference() if kind == "reference" else T.oracle()
    rows = 128
    res = T.aligned(rows * TXQ_W, np.int16).reshape(rows, TXQ_W)
    res[:] = txq_residual(3, 1)[:rows]
    _CPU["res"] = res
    _CPU["out"] = [T.aligned(rows * TXQ_W, np.int32) for _ in range(3)] + [T.aligned(rows * TXQ_W // 16, np.uint16), T.aligned(rows * TXQ_W // 16, np.uint64)]


def _txq_cpu_task(job):
    """One 1920x128 residual strip of one (size, type) through the reference's transform + quantiser on one core."""
    w, h, t, q10 = job
    lib, res, out = _CPU["lib"], _CPU["res"], _CPU["out"]
    fn = lib.ref_txq_plane if _CPU["kind"] == "reference" else lib.meo_txq_plane
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_void_p] * 6
    q = np.ascontiguousarray(q10, dtype=np.int16)
    n = 0
    for _ in range(TXQ_CPU_REPS):
        n += fn(res.ctypes.data, TXQ_W, TXQ_W, res.shape[0], w, h, t, q.ctypes.data, *[o.ctypes.data for o in out])
    return n * w * h


class TxqCpuArm:
    def __init__(self):
        import multiprocessing as mp
        self.kind = "reference" if cpu_reference_available() else "port"
        self.cores = len(os.sched_getaffinity(0))
        self.pool = mp.get_context("fork").Pool(self.cores, initializer=_txq_cpu_init, initargs=(self.kind,))

    def run(self, case, q10, strips):
        t0 = time.perf_counter()
        pix = sum(self.pool.map(_txq_cpu_task, [case + (list(q10),)] * strips, chunksize=1))
        return pix / (time.perf_counter() - t0)

    def close(self):
        self.pool.close()
        self.pool.join()


def txq_quant_params():
    """Luma quantiser of qindex 124, 8-bit: from the golden file (the reference's own av1_build_quantizer tables)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "txq_golden.npz"))
    for c, q in zip(g["cases"], g["q10"]):
        if int(c[3]) == 8 and int(c[4]) == 124:
            return [int(v) for v in q]
    return [int(v) for v in g["q10"][0]]


def run_txq(args, rank, world, local_rank):
    """Sweep of svttxq_fwd_txfm_quant over TXQ_PLANES stacked 1080p residual planes per launch. One JSON line:
    value = pixels/s over the whole sweep (resident), e2e through host buffers, per-case GB/s against the HBM peak."""
    q10 = txq_quant_params()
    cases = txq_cases()
    if args.impl == "reference":
        if rank != 0:
            return
        arm = TxqCpuArm()
        strips = 4 * arm.cores
        per = {}
        for _ in range(max(1, args.warmup // 3)):
            arm.run(cases[0], q10, arm.cores)
        t0 = time.perf_counter()
        pix_total = 0
        for _ in range(args.steps):
            for c in cases:
                r = arm.run(c, q10, strips)
                per["%dx%d/%d" % c] = r / 1e6
                pix_total += strips * TXQ_CPU_REPS * 128 * TXQ_W
        dt = time.perf_counter() - t0
        arm.close()
        v = pix_total / dt / 1e6
        sample = f"{strips} x {TXQ_CPU_REPS} strips of 1920x128 residual per case and step, {len(cases)} cases, {arm.cores} worker processes"
        print(json.dumps({"impl": "reference", "metric": "txq_mpixels_per_sec", "value": v, "unit": "Mpixel/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": txq_config(world),
                          "cpu_baseline": {"value": v, "unit": "Mpixel/s", "cores": arm.cores, "kind": arm.kind, "sample": sample},
                          "e2e": {"value": v, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "per_case_mpixels_per_sec": per, "gpu_launches": 0}), flush=True)
        return

    import torch
    pkg = importlib.import_module("svt-av1-1_b200")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        arm = TxqCpuArm()
        arm.run(cases[0], q10, arm.cores)
        per, t0, pix = {}, time.perf_counter(), 0
        strips = 4 * arm.cores
        for c in cases:
            per["%dx%d/%d" % c] = arm.run(c, q10, strips) / 1e6
            pix += strips * TXQ_CPU_REPS * 128 * TXQ_W
        dt = time.perf_counter() - t0
        arm.close()
        cpu = {"value": pix / dt / 1e6, "unit": "Mpixel/s", "cores": arm.cores, "kind": arm.kind,
               "sample": f"{strips} x {TXQ_CPU_REPS} strips of 1920x128 residual per case, {len(cases)} cases, {arm.cores} worker processes, {dt:.1f} s wall",
               "per_case": per}
    torch.cuda.set_device(local_rank)
    res = txq_residual(11 + rank, TXQ_PLANES)
    host = torch.from_numpy(res).pin_memory()
    dev = host.cuda()
    Hs = TXQ_PLANES * TXQ_H
    npix = Hs * TXQ_W
    d_out = [torch.empty(npix, dtype=torch.int32, device="cuda") for _ in range(3)]
    d_eob = torch.empty(npix // 16, dtype=torch.int16, device="cuda")
    d_en = torch.empty(npix // 16, dtype=torch.int64, device="cuda")
    h_out = [torch.empty(npix, dtype=torch.int32).pin_memory() for _ in range(3)]
    h_eob = torch.empty(npix // 16, dtype=torch.int16).pin_memory()
    h_en = torch.empty(npix // 16, dtype=torch.int64).pin_memory()
    h_lv = torch.empty(npix, dtype=torch.int16).pin_memory()              # compact return path: levels, offsets (eob reuses h_eob)
    h_off = torch.empty(npix // 16 + 1, dtype=torch.int32).pin_memory()
    ctx = pkg.TxqContext(local_rank)
    st = torch.cuda.ExternalStream(ctx.stream())
    quant = pkg.SvtTxqQuant.from_q10(q10)

    def launch(c, energy=True):
        ctx.fwd_txfm_quant_device(dev.data_ptr(), TXQ_W, TXQ_W, Hs, c[0], c[1], c[2], quant, d_out[0].data_ptr(), d_out[1].data_ptr(),
                                  d_out[2].data_ptr(), d_eob.data_ptr(), d_en.data_ptr() if energy else None)

    def host_call(c, everything=True):
        """everything: all five outputs (what Av1EstimateTransform + the quantiser produce); else the subset EncDec consumes
        after quantisation (quantised levels for entropy coding, dequantised for reconstruction, eob)."""
        rc = ctx.lib.svttxq_fwd_txfm_quant(ctx._h, host.data_ptr(), TXQ_W, TXQ_W, Hs, c[0], c[1], c[2], C.byref(quant),
                                           h_out[0].data_ptr() if everything else None, h_out[1].data_ptr(), h_out[2].data_ptr(),
                                           h_eob.data_ptr(), h_en.data_ptr() if everything else None)
        assert rc == 0

    compact_levels = {}

    def compact_call(c):
        """the eob-bounded return path: eob + offsets + the eob leading levels of every block in scan order (int16)"""
        n = C.c_size_t(0)
        rc = ctx.lib.svttxq_fwd_txfm_quant_compact(ctx._h, host.data_ptr(), TXQ_W, TXQ_W, Hs, c[0], c[1], c[2], C.byref(quant),
                                                   h_eob.data_ptr(), h_off.data_ptr(), h_lv.data_ptr(), npix, C.byref(n))
        assert rc == 0
        compact_levels[c] = n.value

    def sync_all():
        ctx.synchronize()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    for _ in range(args.warmup):
        for c in cases:
            launch(c)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    per_ms = {}
    sync_all()
    l0 = ctx.kernel_launches()
    e_all0, e_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    evs = []
    e_all0.record(st)
    for _ in range(args.steps):
        for c in cases:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            launch(c)
            b.record(st)
            evs.append((c, a, b))
    e_all1.record(st)
    sync_all()
    launches = ctx.kernel_launches() - l0
    ms = e_all0.elapsed_time(e_all1)
    for c, a, b in evs:
        per_ms.setdefault(c, []).append(a.elapsed_time(b))
    clocks = sampler.stop() if sampler else None
    # outside the sweep: the 64-point sizes as EncDec calls them (no three_quad_energy wanted -> pruned kernels)
    no_energy_ms = {}
    for c in cases:
        if 64 in c[:2]:
            ts = []
            for _ in range(args.warmup + args.steps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                launch(c, energy=False)
                b.record(st)
                ts.append((a, b))
            sync_all()
            no_energy_ms["%dx%d/%d" % c] = float(np.median([a.elapsed_time(b) for a, b in ts[args.warmup:]]))
    # e2e: host buffers through the synchronous C-ABI call, H2D + kernel + D2H per call
    for c in cases[:2]:
        host_call(c)
    sync_all()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 2))
    for _ in range(e2e_steps):
        for c in cases:
            host_call(c)
    sync_all()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        for c in cases:
            host_call(c, everything=False)
    sync_all()
    ms_e2e_sub = (time.perf_counter() - t0) * 1e3
    for c in cases[:2]:
        compact_call(c)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        for c in cases:
            compact_call(c)
    sync_all()
    ms_e2e_compact = (time.perf_counter() - t0) * 1e3
    if dist is not None:
        t = torch.tensor([ms, ms_e2e, ms_e2e_sub, ms_e2e_compact], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e, ms_e2e_sub, ms_e2e_compact = float(t[0].item()), float(t[1].item()), float(t[2].item()), float(t[3].item())
    checksum = int(h_out[1].sum(dtype=torch.int64).item())
    ctx.close()
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        per = {}
        tot_bytes = tot_ms = 0.0
        for c in cases:
            w, h, _ = c
            pw, ph = min(w, 32), min(h, 32)
            nb = npix // (w * h)
            nbytes = 2 * npix + 12 * nb * pw * ph + 10 * nb
            t = float(np.median(per_ms[c]))
            per["%dx%d/%d" % c] = {"ms": t, "GB/s": nbytes / t / 1e6, "frac": nbytes / t / 1e6 / hbm_peak, "Gpixel/s": npix / t / 1e6}
            tot_bytes += nbytes
            tot_ms += t
        worst = min(per, key=lambda k: per[k]["frac"])
        # DRAM bytes per launch: the byte model of every case scaled by the measured / model ratio of its size class in the
        # committed ncu capture (profiles/r02/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one launch)
        traffic = None
        try:
            tc = json.load(open(os.path.join(ROOT, "profiles", "r02", "traffic.json")))["txq_kernel"]["cases"]
            small = np.mean([v["ratio"] for k, v in tc.items() if "64" not in k])
            big = np.mean([v["ratio"] for k, v in tc.items() if "64" in k])
            tb = 0.0
            for (w, h, _) in cases:
                nb = npix // (w * h)
                tb += (2 * npix + 12 * nb * min(w, 32) * min(h, 32) + 10 * nb) * (big if 64 in (w, h) else small)
            traffic = tb / len(cases)
        except Exception:
            pass
        v = world * npix * len(cases) * args.steps / (ms / 1e3) / 1e6
        v_e2e = world * npix * len(cases) * e2e_steps / (ms_e2e / 1e3) / 1e6
        v_compact = world * npix * len(cases) * e2e_steps / (ms_e2e_compact / 1e3) / 1e6
        compact_bytes = sum(2 * compact_levels[c] + 6 * (npix // (c[0] * c[1])) + 4 for c in cases)
        pw_bytes = sum(12 * (npix // (w * h)) * min(w, 32) * min(h, 32) + 10 * (npix // (w * h)) for (w, h, _) in cases)
        line = {"metric": "txq_mpixels_per_sec", "value": v, "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                "data": "synthetic", "config": txq_config(world),
                "e2e": {"value": v_compact, "unit": "Mpixel/s", "h2d_bytes_per_step": 2 * npix * len(cases), "d2h_bytes_per_step": compact_bytes,
                        "ms_per_step": ms_e2e_compact / e2e_steps, "call": "svttxq_fwd_txfm_quant_compact",
                        "outputs": "per block: eob (u16), level offset (u32) and the eob leading quantised levels in scan order (i16) -- "
                                   "what the entropy coder reads; dqcoeff is rebuilt by the caller (EbFullLoop.c:596-662)",
                        "kept_levels_per_pixel": sum(compact_levels.values()) / (npix * len(cases)),
                        "timing": "host wall clock around synchronous calls, pinned host buffers"},
                "e2e_all_outputs": {"value": v_e2e, "unit": "Mpixel/s", "h2d_bytes_per_step": 2 * npix * len(cases), "d2h_bytes_per_step": pw_bytes,
                                    "ms_per_step": ms_e2e / e2e_steps, "call": "svttxq_fwd_txfm_quant (coeff + qcoeff + dqcoeff as int32, eob, energy)"},
                "e2e_encdec_outputs": {"value": world * npix * len(cases) * e2e_steps / (ms_e2e_sub / 1e3) / 1e6, "unit": "Mpixel/s",
                                       "outputs": "qcoeff + dqcoeff + eob (coeff and energy not copied back)",
                                       "d2h_bytes_per_step": sum(8 * (npix // (w * h)) * min(w, 32) * min(h, 32) + 2 * (npix // (w * h)) for (w, h, _) in cases)},
                "gpu_launches": int(launches),
                "roofline": {"kernel": "txq_kernel<*>", "bound": "hbm", "achieved": tot_bytes / tot_ms / 1e6, "peak": hbm_peak, "unit": "GB/s",
                             "frac": tot_bytes / tot_ms / 1e6 / hbm_peak, "traffic": traffic,
                             "traffic_source": "average DRAM bytes per launch over the sweep: byte model x measured/model ratio per size class from "
                                               "profiles/r02/traffic.json (ncu --set full: 0.75 up to 32 points, 0.47-0.59 for 64-point sizes; below the "
                                               "model because part of a launch's 200 MB of output is still dirty in the 126 MB L2 when the kernel ends)",
                             "algorithmic_bytes_per_launch": tot_bytes / len(cases),
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                             "bytes_model": "2 B/pixel residual in + 12 B per kept coefficient (coeff, qcoeff, dqcoeff) + 10 B per block (eob, energy)",
                             "worst_case": worst, "worst_frac": per[worst]["frac"]},
                "per_case": per,
                "per_case_without_energy": {k: {"ms": t, "Gpixel/s": npix / t / 1e6, "speedup_vs_with_energy": per[k]["ms"] / t}
                                            for k, t in no_energy_ms.items()},
                "clocks": clocks, "result_checksum": checksum}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def txq_config(n_gpus):
    return {"workload": f"forward transform + quantisation sweep: {len(txq_cases())} (size, type) cases over the 19 AV1 transform sizes "
                        f"(DCT_DCT everywhere, ADST_ADST up to 16 points, IDTX up to 32), each over {TXQ_PLANES} stacked 1920x1088 "
                        "8-bit residual planes per launch, luma quantiser of qindex 124",
            "pixels_per_launch": TXQ_PLANES * TXQ_H * TXQ_W, "parallelism": f"independent planes per GPU x{n_gpus}, no collective",
            "l2_policy": "outputs of one launch (200 MB) exceed L2; consecutive launches use different kernels over the same 33 MB input"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the sharded-mode measurements (strong_1080p / strong_4k keys)")
    ap.add_argument("--search-area", type=int, default=0,
                    help="full-pel search area N x N (1..255; the reference searches MIN(N, 127)) instead of the preset's 64 x 64")
    ap.add_argument("--workload", default="me", choices=["me", "me4k", "me209", "streams", "txq"],
                    help="me: 1080p motion estimation (headline, BASELINE configs[1]); me4k: 3840x2160 10-bit, SB rows sharded over "
                         "the GPUs (configs[2]); txq: transform + quantisation sweep (configs[3])")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "txq":
        run_txq(args, rank, world, local_rank)
        return
    select_workload(args.workload)
    if args.search_area:
        global SEARCH_AREA
        SEARCH_AREA = args.search_area
    if args.impl == "reference":
        run_reference_impl(args, rank)
        return
    run_gpu_impl(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
