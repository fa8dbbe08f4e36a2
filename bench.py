#!/usr/bin/env python
"""bench.py - headline measurement of the Packager texture stage on B200 (driver contract).

Metric (BASELINE.json): BC7+mips Mpix/s.  Workload = BASELINE.json configs[4] (the throughput
sweep: 512 synthetic 2048x2048 albedo textures, full mip chain + BC7 at the reference's
Quality_Normal), sharded per texture with no collective: every rank (= GPU) owns 64 textures per
step, so 8 GPUs process the config's whole 512-texture set in one step ("scaling": "weak").
Mpix = level-0 pixels; every texture additionally carries its 11 smaller levels.

One "step" = one pass of the hot path over the rank's batch:
  value : inputs already resident in HBM, rosy_tex_encode_batch_device() on one stream (one mip launch
          and one BC7 launch for the step's textures; --per-texture: rosy_tex_mips_device /
          rosy_tex_blocks_device per texture),
          timed with CUDA events on that stream (barrier + synchronize both sides, max over ranks)
  e2e   : the same batch through the reference-facing C ABI with HOST buffers
          (rosy_tex_submit_rgba8 + rosy_tex_wait_all: pinned H2D, kernels, D2H of the .dds bytes),
          timed by the host clock around the calls, max over ranks
  roofline     : per-kernel CUDA-event durations taken inside the timed region
  cpu_baseline : the CPU oracle (oracle/, "port" - NVTT itself is closed source and absent) on a
                 bounded sample of the same textures, all host cores, rank 0 at N = 1 only

`--impl reference` times that CPU oracle as the reference arm (the only CPU implementation of the
path that exists here).  The product path never touches oracle/.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "BC7+mips Mpix/s"
UNIT = "Mpix/s"
SM_COUNT = 148
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--textures", type=int, default=64, help="textures per GPU per step")
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--quality", default="normal", choices=["fast", "normal", "high"])
    ap.add_argument("--kind", default="albedo", choices=["albedo", "normal", "metal_rough"])
    ap.add_argument("--format", default="bc7", choices=["bc7", "bc5", "bc4"])
    ap.add_argument("--cpu-sample", type=int, default=2, help="textures in the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the e2e leg")
    ap.add_argument("--files", type=int, default=16, help="image files in the e2e_files leg (0: skip)")
    ap.add_argument("--per-texture", action="store_true",
                    help="device leg: one mip launch per texture (rosy_tex_mips_device) instead of the batched call")
    ap.add_argument("--split-stages", action="store_true",
                    help="BC5 / BC4: mips and blocks as two calls (two kernels timed apart) instead of the fused mip + encode kernel")
    return ap.parse_args()


# ------------------------------------------------------------------------------------ inputs
def make_textures(kind, size, count, rank):
    """`count` distinct synthetic textures (SURVEY.md 8d generators).  A few seeded bases are
    generated with the full noise stack; the rest are cheap re-arrangements of them (odd rolls,
    flips, transposes, channel rotations), which keeps start-up at seconds while every texture
    still has different 4x4 blocks."""
    from rosy_b200 import synth

    nbase = min(count, 4)
    bases = [synth.make(kind, size, size, 16 * rank + i) for i in range(nbase)]
    out = []
    for i in range(count):
        b = bases[i % nbase]
        v = i // nbase
        if v:
            b = np.roll(b, (37 * v + 1, 53 * v + 3), axis=(0, 1))
            if v & 1:
                b = b[::-1]
            if v & 2:
                b = b[:, ::-1]
            if v & 4:
                b = b.transpose(1, 0, 2)
            if kind == "albedo" and v >= 8:
                b = b[..., [(0 + v) % 3, (1 + v) % 3, (2 + v) % 3, 3]]
        out.append(np.ascontiguousarray(b))
    return out


# ------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""

    REASONS = {
        0x8: "hw_slowdown",
        0x40: "hw_thermal_slowdown",
        0x20: "sw_thermal_slowdown",
        0x4: "sw_power_cap",
        0x80: "hw_power_brake_slowdown",
    }

    def __init__(self, torch, device_index):
        self.samples, self.reasons, self.sm_max = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv is not None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": 0}
        return {
            "sm_mhz": float(np.median(self.samples)),
            "sm_max_mhz": self.sm_max,
            "reasons": sorted(self.reasons),
            "samples": len(self.samples),
        }


# ------------------------------------------------------------------------------------ cpu legs
def cpu_oracle_run(textures, kind, fmt, quality, threads=0):
    """One pass of the CPU oracle over `textures`; returns seconds."""
    import oracle as O

    okind = O.KIND_NORMAL if kind == "normal" else O.KIND_SRGB
    t0 = time.perf_counter()
    for img in textures:
        O.generate_texture(img, okind, fmt, quality, threads=threads)
    return time.perf_counter() - t0


def run_reference_arm(args, rank):
    """--impl reference: the CPU implementation of the path on this box's host cores.

    NVTT 3.2.5 (what the reference links, src/premake5.lua:123) is a closed-source Windows DLL that
    is neither in /root/reference nor in this image, so the reference's own CPU code for the path
    cannot be built; the timed CPU arm is the oracle port (oracle/, kind = "port")."""
    if rank != 0:
        return
    import oracle as O

    fmt = {"bc7": 0, "bc5": 1, "bc4": 2}[args.format]
    q = {"fast": 0, "normal": 1, "high": 2}[args.quality]
    sample = make_textures(args.kind, args.size, 1, 0)
    cores = O.max_threads()
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_oracle_run(sample, args.kind, fmt, q)
    times = [cpu_oracle_run(sample, args.kind, fmt, q) for _ in range(args.steps)]
    sec = float(np.mean(times))
    mpix = len(sample) * args.size * args.size / 1e6
    value = mpix / sec
    sec1 = cpu_oracle_run(sample, args.kind, fmt, q, threads=1)
    # the same `config` as the GPU arm's line, key for key (the driver compares them); what this arm's bounded step
    # covers is said in cpu_baseline.sample
    cfg = workload_config(args, args.textures)
    line = {
        "impl": "reference",
        "metric": METRIC,
        "value": value,
        "unit": UNIT,
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": sec * 1e3,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "u8+fp32",
        "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {
            "value": value,
            "unit": UNIT,
            "cores": cores,
            "value_1t": mpix / sec1,
            "kind": "port",
            "sample": f"each timed step of this arm is 1 of the workload's {args.textures} textures per GPU (bounded CPU sample; the "
            f"metric is per pixel, so the value is comparable): 1 x {args.size}x{args.size} {args.kind}, full mips + {args.format.upper()} {args.quality}; "
            "mip rows and blocks both shared by all host threads; value_1t = the same texture on 1 thread",
        },
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, ntex, note=None):
    c = {
        "workload": f"configs[4] throughput sweep: {args.size}x{args.size} synthetic {args.kind}, full mip chain + "
        f"{args.format.upper()} at quality {args.quality}, sharded per texture "
        f"({args.textures} textures per GPU per step; 8 GPUs = the config's 512)",
        "textures_per_gpu": ntex,
        "size": args.size,
        "format": args.format,
        "quality": args.quality,
        "parallelism": f"texture-sharded x{args.gpus}, no collective",
        "l2": "inputs per step exceed the 126 MB L2 (no flush needed)"
        if ntex * args.size * args.size * 4 > L2_BYTES
        else "L2 flushed between steps",
    }
    if note:
        c["note"] = note
    return c


# ------------------------------------------------------------------------------------ main
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist

    from rosy_b200 import texture as T

    T.lib()  # fails loudly when the CUDA extension is missing
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # The path has no exchange step: the only multi-rank traffic of this script is the barrier and
        # the MAX of the times, which go over gloo on the host.  NCCL is not initialised anywhere (north_star).
        dist.init_process_group("gloo")

    fmt = {"bc7": T.FORMAT_BC7, "bc5": T.FORMAT_BC5, "bc4": T.FORMAT_BC4}[args.format]
    q = {"fast": 0, "normal": 1, "high": 2}[args.quality]
    kind = {"albedo": T.KIND_COLOR, "normal": T.KIND_NORMAL_MAP, "metal_rough": T.KIND_METALLIC_ROUGHNESS}[args.kind]
    T.init(devices=[local_rank], quality=q)

    n, size = args.textures, args.size
    textures = make_textures(args.kind, size, n, rank)
    d = T.desc(size, size, kind, fmt, q, -1)
    in_bytes = size * size * 4
    dds_bytes = T.dds_size(d)
    blk_bytes = dds_bytes - T.DDS_HEADER_BYTES
    chain_bytes = T.chain_size(size, size)
    mpix_step = n * size * size / 1e6

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident leg: `value` + per-kernel durations --------------------
    d_in = torch.empty((n, in_bytes), dtype=torch.uint8, device=dev)
    for i, img in enumerate(textures):
        d_in[i].copy_(torch.from_numpy(img.reshape(-1)))
    # one chain / scratch buffer per texture: the step is ONE batched call (one mip launch for the
    # whole batch, then one BC7 launch for the whole batch)
    scratch_bytes = T.device_scratch_size(d)
    chain_pitch = (chain_bytes + 255) // 256 * 256  # device pointers must be 16-byte aligned
    d_chain = torch.empty((n, chain_pitch), dtype=torch.uint8, device=dev)
    d_scratch = torch.zeros((n, scratch_bytes), dtype=torch.uint8, device=dev)
    d_blocks = torch.empty((n, (blk_bytes + 255) // 256 * 256), dtype=torch.uint8, device=dev)
    d_batch = torch.zeros(T.batch_scratch_size(n), dtype=torch.uint8, device=dev)
    batch = T.make_batch([(d, d_in[i].data_ptr(), d_chain[i].data_ptr(), d_scratch[i].data_ptr(),
                           d_blocks[i].data_ptr()) for i in range(n)])
    flush = None
    if n * in_bytes <= L2_BYTES:
        flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    fused = fmt != T.FORMAT_BC7 and not args.split_stages and not args.per_texture and os.environ.get("ROSY_TEX_FUSE_ENCODE", "1") != "0"

    def step(events=None):
        if flush is not None:
            flush.fill_(1)
        if args.per_texture:
            for i in range(n):
                if events is not None:
                    ea, eb, ec = events[i]
                    ea.record()
                T.mips_device(d, d_in[i].data_ptr(), d_chain[i].data_ptr(), d_scratch[i].data_ptr(), stream)
                if events is not None:
                    eb.record()
                T.blocks_device(d, d_in[i].data_ptr(), d_chain[i].data_ptr(), d_blocks[i].data_ptr(), stream)
                if events is not None:
                    ec.record()
            return
        if events is not None:
            ea, eb, ec = events[0]
            ea.record()
        if fused:
            # BC5 / BC4: ONE call for both stages, so that the fused mip + encode kernel runs (levels 0..4 are encoded
            # from the tiles the mip kernel holds; a small bc45_encode_kernel launch takes levels 5..)
            T.encode_batch_device(batch, d_batch.data_ptr(), stream, True, True)
            if events is not None:
                eb.record()
                ec.record()
            return
        T.encode_batch_device(batch, d_batch.data_ptr(), stream, True, False)
        if events is not None:
            eb.record()
        T.encode_batch_device(batch, d_batch.data_ptr(), stream, False, True)
        if events is not None:
            ec.record()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    clocks = ClockSampler(torch, local_rank)
    kernel_events = [
        [tuple(torch.cuda.Event(enable_timing=True) for _ in range(3)) for _ in range(n)] for _ in range(args.steps)
    ]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = T.kernel_launches()
    clocks.start()
    barrier()
    e0.record()
    for k in range(args.steps):
        step(kernel_events[k])
    e1.record()
    barrier()
    clocks.stop()
    launches = T.kernel_launches() - launches0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_step = ms_total / args.steps
    value = world * mpix_step / (ms_step * 1e-3)
    # per-launch durations: the batched step has ONE mip launch (all n textures) and ONE encoder launch
    used = [evs if args.per_texture else evs[:1] for evs in kernel_events]
    mip_ms = float(np.mean([ea.elapsed_time(eb) for evs in used for (ea, eb, ec) in evs]))
    enc_ms = float(np.mean([eb.elapsed_time(ec) for evs in used for (ea, eb, ec) in evs]))
    mip_tex_per_launch = 1 if args.per_texture else n
    # the batched call encodes every texture of the step in ONE launch (per format / BC7 quality tier)
    enc_tex_per_launch = 1 if args.per_texture else n
    clk = clocks.summary()

    # ---------------- e2e leg: host buffers through the C ABI work queue --------------------
    e2e = None
    if not args.no_e2e:
        L = T.lib()
        host_in = [L.rosy_tex_alloc_host(in_bytes) for _ in range(n)]
        host_out = [L.rosy_tex_alloc_host(dds_bytes) for _ in range(n)]
        if not all(host_in) or not all(host_out):
            raise SystemExit("bench.py: pinned host allocation failed")
        for p, img in zip(host_in, textures):
            C.memmove(p, img.ctypes.data, in_bytes)
        lens = [C.c_size_t() for _ in range(n)]

        def e2e_step():
            for i in range(n):
                rc = L.rosy_tex_submit_rgba8(C.byref(d), host_in[i], host_out[i], dds_bytes, C.byref(lens[i]))
                if rc:
                    raise SystemExit(f"rosy_tex_submit_rgba8 failed: {L.rosy_tex_last_error().decode()}")
            rc = L.rosy_tex_wait_all()
            if rc:
                raise SystemExit(f"rosy_tex_wait_all failed: {L.rosy_tex_last_error().decode()}")
            # device->host read of the step's result: the first bytes of the last .dds (already in host memory)
            return C.string_at(host_out[n - 1], 4)

        for _ in range(max(1, min(args.warmup, 2))):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            magic = e2e_step()
        torch.cuda.synchronize()
        sec = time.perf_counter() - t0
        barrier()
        assert magic == b"DDS ", magic
        # the e2e output must be the same bytes as the device-resident leg produced
        got = np.ctypeslib.as_array(C.cast(host_out[0], C.POINTER(C.c_uint8)), shape=(dds_bytes,))
        assert (got[T.DDS_HEADER_BYTES:] == d_blocks[0, :blk_bytes].cpu().numpy()).all(), "e2e and device legs disagree"
        sec = max_over_ranks(sec)
        e2e = {
            "value": world * mpix_step * args.steps / sec,
            "unit": UNIT,
            "h2d_bytes_per_step": n * in_bytes,
            "d2h_bytes_per_step": n * blk_bytes,
            "ms_per_step": sec * 1e3 / args.steps,
            "api": "rosy_tex_submit_rgba8 + rosy_tex_wait_all (pinned host buffers, 2 slots per device)",
        }
        for p in host_in + host_out:
            L.rosy_tex_free_host(p)

        # ---- the host-feed ceiling beside it (SURVEY 8d "Scaling"): bare cudaMemcpyAsync, no kernels, the same byte mix per
        # texture (16 MiB up, 5.33 MiB down at 2048^2), pinned buffers, two streams, every rank at the same time.  What e2e
        # could reach if the copies were all there is; tools/h2d_ceiling.cu is the same measurement without Python.
        depth = 4
        h_up = [torch.empty(in_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(depth)]
        h_dn = [torch.empty(blk_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(depth)]
        d_up = [torch.empty(in_bytes, dtype=torch.uint8, device=dev) for _ in range(depth)]
        d_dn = [torch.zeros(blk_bytes, dtype=torch.uint8, device=dev) for _ in range(depth)]
        s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def copy_round():
            with torch.cuda.stream(s_up):
                for k in range(depth):
                    d_up[k].copy_(h_up[k], non_blocking=True)
            with torch.cuda.stream(s_dn):
                for k in range(depth):
                    h_dn[k].copy_(d_dn[k], non_blocking=True)
            s_up.synchronize()
            s_dn.synchronize()

        copy_round()
        barrier()
        t0 = time.perf_counter()
        rounds = 0
        while time.perf_counter() - t0 < 1.0:
            copy_round()
            rounds += 1
        csec = time.perf_counter() - t0
        barrier()
        up_gbs = rounds * depth * in_bytes / csec / 1e9
        dn_gbs = rounds * depth * blk_bytes / csec / 1e9
        if world > 1:
            tot = torch.tensor([up_gbs, dn_gbs], dtype=torch.float64)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
            up_gbs, dn_gbs = float(tot[0]), float(tot[1])
        ceiling_mpix = up_gbs * 1e9 / 4 / 1e6  # 4 bytes of level 0 per pixel go up
        e2e["host_ceiling_gbs"] = up_gbs
        e2e["host_ceiling"] = {
            "h2d_gbs": up_gbs,
            "d2h_gbs": dn_gbs,
            "mpix_s": ceiling_mpix,
            "e2e_frac_of_ceiling": e2e["value"] / ceiling_mpix,
            "how": f"bare cudaMemcpyAsync from / to pinned memory on all {world} rank(s) at once, {in_bytes} B up + {blk_bytes} B down "
            "per texture, no kernels, 1 s; sum over ranks",
        }
        del h_up, h_dn, d_up, d_dn

    # ---------------- e2e_files leg: image files -> .dds files through rosy_tex_submit -----------
    # (the reference interface itself: generate_*_texture(path), Packager.cpp:216,263; decode + write included)
    e2e_files = None
    if rank == 0 and world == 1 and args.files > 0 and not args.no_e2e:
        e2e_files = files_leg(T, textures[: min(args.files, n)], kind, size)

    # ---------------- roofline of the dominant kernel ----------------------------------------
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    npx = size * size
    # algorithmic bytes per launch: mip kernel 4N + 4N/3 (+4N for normal maps, whose level 0
    # is rewritten); encoders read 4 B/px and write 1 (BC7/BC5) or 0.5 (BC4) B/px over 4N/3 px
    chain_px = chain_bytes // 4
    mip_bytes = (4 * npx + 4 * (chain_px - npx) + (4 * npx if kind == T.KIND_NORMAL_MAP else 0)) * mip_tex_per_launch
    enc_bytes = (4 * chain_px + blk_bytes) * enc_tex_per_launch
    if fused:
        # SURVEY 8(d) fused floor: level 0 read once, every level's blocks written once
        mip_bytes = (4 * npx + blk_bytes) * mip_tex_per_launch
    mip_gbs = mip_bytes / (mip_ms * 1e-3) / 1e9
    enc_gbs = enc_bytes / (enc_ms * 1e-3) / 1e9 if enc_ms > 0 else 0.0
    prof = load_profile_counts(args)
    sm_mhz = clk["sm_mhz"] or float(peaks.get("sm_max_mhz", 1965.0))
    issue_peak = SM_COUNT * 4 * sm_mhz * 1e6 / 1e9  # G warp-instructions / s at the clock seen under load
    enc_kernel = "bc7_encode_kernel" if fmt == T.FORMAT_BC7 else "bc45_encode_kernel"
    if fmt == T.FORMAT_BC7:
        roofline = {
            "kernel": enc_kernel,
            "bound": "issue",
            "achieved": None,
            "peak": issue_peak,
            "unit": "Gwarp-inst/s",
            "frac": None,
            "traffic": (prof["bc7_dram_bytes_per_texture"] * enc_tex_per_launch if prof.get("bc7_dram_bytes_per_texture") else None),
            "ms_per_launch": enc_ms,
            "textures_per_launch": enc_tex_per_launch,
            "share_of_step": enc_ms * (n // enc_tex_per_launch) / ms_step,
            "hbm": {"achieved": enc_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": enc_gbs / hbm_peak,
                    "algorithmic_bytes_per_launch": enc_bytes, "peak_source": peak_src},
            "note": "BC7 is issue-slot bound (north_star): achieved = ncu smsp__inst_executed.sum per launch "
            "(profiles/) / live CUDA-event duration; peak = 148 SMs x 4 schedulers x SM clock under load",
        }
        if prof.get("bc7_warp_inst_per_texture"):
            a = prof["bc7_warp_inst_per_texture"] * enc_tex_per_launch / (enc_ms * 1e-3) / 1e9
            roofline["achieved"] = a
            roofline["frac"] = a / issue_peak
            roofline["inst_source"] = prof.get("source")
    elif fused:
        roofline = {
            "kernel": "mip_fused_kernel<ENC> (mip chain + BC5/BC4 blocks of levels 0..4) + bc45_encode_kernel (levels 5..)",
            "bound": "hbm", "achieved": mip_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": mip_gbs / hbm_peak, "traffic": None,
            "ms_per_launch": mip_ms, "share_of_step": mip_ms / ms_step, "algorithmic_bytes_per_launch": mip_bytes,
            "textures_per_launch": mip_tex_per_launch, "peak_source": peak_src,
            "note": "algorithmic bytes = SURVEY 8(d)'s fused floor: 4 N read + the block stream written; the separate stages "
            "(--split-stages) move 4 N + 4 N/3 (+ 4 N for normal maps) and then 4 (4 N/3) + blocks",
        }
    else:
        roofline = {
            "kernel": enc_kernel, "bound": "hbm", "achieved": enc_gbs, "peak": hbm_peak, "unit": "GB/s",
            "frac": enc_gbs / hbm_peak, "traffic": None, "ms_per_launch": enc_ms,
            "share_of_step": enc_ms * (n // enc_tex_per_launch) / ms_step, "algorithmic_bytes_per_launch": enc_bytes,
            "peak_source": peak_src,
        }
    roofline_mip = None if fused else {
        "kernel": "mip_fused_kernel", "bound": "hbm", "achieved": mip_gbs, "peak": hbm_peak, "unit": "GB/s",
        "frac": mip_gbs / hbm_peak, "traffic": (prof["mip_dram_bytes_per_texture"] * mip_tex_per_launch if prof.get("mip_dram_bytes_per_texture") else None),
        "ms_per_launch": mip_ms,
        "share_of_step": mip_ms * (n // mip_tex_per_launch) / ms_step, "algorithmic_bytes_per_launch": mip_bytes,
        "textures_per_launch": mip_tex_per_launch, "peak_source": peak_src,
    }

    # ---------------- cpu_baseline (rank 0, N = 1 only) ----------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as O

        sample = textures[: max(1, args.cpu_sample)]
        sec = cpu_oracle_run(sample, args.kind, fmt, q)  # every core: mip rows and blocks both threaded
        sec1 = cpu_oracle_run(sample[:1], args.kind, fmt, q, threads=1)
        cpu = {
            "value": len(sample) * npx / 1e6 / sec,
            "unit": UNIT,
            "cores": O.max_threads(),
            "value_1t": npx / 1e6 / sec1,
            "kind": "port",
            "sample": f"{len(sample)} of the step's {n} textures ({size}x{size} {args.kind}, full mips + "
            f"{args.format.upper()} {args.quality}), {sec:.1f} s on all cores; value_1t: 1 texture on 1 thread, {sec1:.1f} s; "
            "oracle/ built -O3 -march=x86-64-v3 (the .so travels between hosts, so not -march=native)",
        }

    if rank == 0:
        line = {
            "metric": METRIC,
            "value": value,
            "unit": UNIT,
            "n_gpus": world,
            "steps": args.steps,
            "warmup": max(3, args.warmup),
            "ms_per_step": ms_step,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "u8+fp32",
            "data": "synthetic",
            "config": workload_config(args, n),
            "e2e": e2e,
            "e2e_files": e2e_files,
            "gpu_launches": int(launches),
            "roofline": roofline,
            "roofline_mip": roofline_mip,
            "cpu_baseline": cpu,
            "clocks": clk,
            "mpix_all_levels_per_step": n * chain_px / 1e6,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    T.shutdown()


def files_leg(T, images, kind, size):
    """PNG / JPEG files on disk -> .dds files next to them (decode threads -> pinned staging -> device slots ->
    writer threads).  Decode-bound on the host; reported beside `e2e`, whose inputs are already decoded."""
    import shutil
    import tempfile

    from PIL import Image

    # a memory-backed directory when there is one: the leg measures decode -> device -> .dds bytes, not the box's overlay
    # disk (268 MB of .dds per run; the container's /tmp sustained ~340 MB/s, which capped the leg at 255 Mpix/s)
    base = None
    try:
        st = os.statvfs("/dev/shm")
        if st.f_bavail * st.f_frsize > (2 << 30):
            base = "/dev/shm"
    except OSError:
        pass
    d = tempfile.mkdtemp(prefix="rosy_bench_files_", dir=base)
    try:
        paths = []
        for i, img in enumerate(images):
            if i % 2:
                p = os.path.join(d, f"t{i}.jpg")
                Image.fromarray(img[..., :3]).save(p, quality=90)
            else:
                p = os.path.join(d, f"t{i}.png")
                Image.fromarray(img).save(p, compress_level=1)
            paths.append(p)
        # three copies of every file: more files than host cores, so that the figure is the pipeline's throughput and
        # not the latency of its slowest single decode
        for p in list(paths):
            for k in (1, 2):
                q = os.path.join(d, f"c{k}_" + os.path.basename(p))
                shutil.copyfile(p, q)
                paths.append(q)
        in_bytes = sum(os.path.getsize(p) for p in paths)
        T.submit_file(paths[0], kind)  # warm-up: I/O threads, pinned pool
        T.wait_all()
        t0 = time.perf_counter()
        for p in paths:
            T.submit_file(p, kind)
        T.wait_all()
        sec = time.perf_counter() - t0
        out_bytes = sum(os.path.getsize(os.path.splitext(p)[0] + ".dds") for p in paths)
        return {"value": len(paths) * size * size / 1e6 / sec, "unit": UNIT, "files": len(paths),
                "formats": "PNG / JPEG q90, half each", "input_file_bytes": in_bytes, "output_file_bytes": out_bytes,
                "seconds": sec, "directory": base or tempfile.gettempdir(), "io_threads": int(os.environ.get("ROSY_TEX_IO_THREADS", os.cpu_count() or 1)),
                "api": "rosy_tex_submit(path) + rosy_tex_wait_all: file decode and .dds write inside the timed region"}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def load_profile_counts(args):
    """ncu-derived per-launch counts for the bench workload (profiles/bench_counts.json), valid only
    for the kernel source they were captured from."""
    path = os.path.join(ROOT, "profiles", "bench_counts.json")
    try:
        with open(path) as f:
            doc = json.load(f)
    except Exception:
        return {}
    key = f"{args.kind}-{args.size}-{args.format}-{args.quality}"
    ent = doc.get(key)
    if not ent:
        return {}
    src = os.path.join(ROOT, "rosy_b200", "csrc", "bc7_encode.cu")
    with open(src, "rb") as f:
        sha = hashlib.sha256(f.read()).hexdigest()[:16]
    if ent.get("bc7_source_sha") and ent["bc7_source_sha"] != sha:
        ent = dict(ent)
        ent.pop("bc7_warp_inst_per_texture", None)  # stale: counted on another version of the kernel
    return ent


if __name__ == "__main__":
    main()
