#!/usr/bin/env python
"""Measure every configuration of BASELINE.json on one B200 (bench.py covers configs[4], the one
the metric is quoted on; this tool fills in the table of DESIGN.md section 6 for the others).

    python tools/run_configs.py [--configs 0,1,2,3,4] [--psnr]  > gpurun_out/configs.jsonl

Per configuration one JSON line: device-resident throughput (one batched call, CUDA events), end to
end throughput through rosy_tex_submit_rgba8 / wait_all from pinned host buffers, and - with --psnr,
on the FULL image of one texture per kind and size - the PSNR of the GPU output next to the CPU oracle's, per level and
chain-weighted (test infrastructure, used here only as the checker).
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # make_textures (cheap distinct textures from a few seeded bases)


def texture_set(cfg):
    """-> list of (kind_name, size, format, quality) for BASELINE.json configs[cfg]"""
    if cfg == 0:
        return "configs[0]: 1 x 1024^2 albedo, BC7 normal", [("albedo", 1024, "bc7", 1)]
    if cfg == 1:
        return "configs[1]: 1 x 4096^2 normal map, BC5 (north_star) ", [("normal", 4096, "bc5", 1)]
    if cfg == 11:
        return "configs[1] with the reference's real format: 1 x 4096^2 normal map, BC7 normal", [("normal", 4096, "bc7", 1)]
    if cfg == 2:
        # 70 textures = 24 albedo + 24 normal + 22 metal-rough; 30 x 1024^2, 30 x 2048^2, 10 x 4096^2
        kinds = ["albedo"] * 24 + ["normal"] * 24 + ["metal_rough"] * 22
        sizes = [1024] * 30 + [2048] * 30 + [4096] * 10
        rng = np.random.RandomState(0x524F5359 & 0x7FFFFFFF)
        rng.shuffle(sizes)
        return ("configs[2]: Sponza-scale set, 70 textures 1K-4K mixed albedo/normal/metal-rough, BC7 normal for all "
                "(the reference's format)"), [(k, s, "bc7", 1) for k, s in zip(kinds, sizes)]
    if cfg == 3:
        return ("configs[3] on ONE GPU: 16 x 8192^2 (8 albedo + 8 normal), BC7 normal",
                [("albedo", 8192, "bc7", 1)] * 8 + [("normal", 8192, "bc7", 1)] * 8)
    if cfg in (40, 41, 42):
        q = cfg - 40
        return (f"configs[4] slice: 64 x 2048^2 albedo, BC7 {['fast', 'normal', 'high'][q]}", [("albedo", 2048, "bc7", q)] * 64)
    raise SystemExit(f"unknown config {cfg}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="0,1,11,2,3,40,41,42")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--psnr", action="store_true")
    args = ap.parse_args()
    import torch

    from rosy_b200 import texture as T

    T.lib()
    dev = torch.device("cuda", 0)
    T.init(devices=[0], quality=1)
    KIND = {"albedo": T.KIND_COLOR, "normal": T.KIND_NORMAL_MAP, "metal_rough": T.KIND_METALLIC_ROUGHNESS}
    FMT = {"bc7": T.FORMAT_BC7, "bc5": T.FORMAT_BC5, "bc4": T.FORMAT_BC4}
    L = T.lib()
    for cfg in [int(x) for x in args.configs.split(",")]:
        name, spec = texture_set(cfg)
        # distinct textures per (kind, size) group, generated the way bench.py does
        groups = {}
        for k, s, f, q in spec:
            groups.setdefault((k, s), 0)
            groups[(k, s)] += 1
        pools = {key: bench.make_textures(key[0], key[1], cnt, 0) for key, cnt in groups.items()}
        taken = {key: 0 for key in groups}
        texs = []
        for k, s, f, q in spec:
            img = pools[(k, s)][taken[(k, s)]]
            taken[(k, s)] += 1
            texs.append((img, T.desc(s, s, KIND[k], FMT[f], q, -1), k))
        mpix = sum(img.shape[0] * img.shape[1] for img, _, _ in texs) / 1e6
        # ---- device-resident: one batched call ----
        keep, entries = [], []
        for img, d, _ in texs:
            t_in = torch.from_numpy(img.reshape(-1)).to(dev)
            t_chain = torch.empty((T.chain_size(img.shape[1], img.shape[0]) + 255) // 256 * 256, dtype=torch.uint8, device=dev)
            t_scr = torch.zeros(T.device_scratch_size(d), dtype=torch.uint8, device=dev)
            t_blk = torch.empty((T.dds_size(d) - T.DDS_HEADER_BYTES + 255) // 256 * 256, dtype=torch.uint8, device=dev)
            keep.append((t_in, t_chain, t_scr, t_blk))
            entries.append((d, t_in.data_ptr(), t_chain.data_ptr(), t_scr.data_ptr(), t_blk.data_ptr()))
        batch = T.make_batch(entries)
        t_batch = torch.zeros(T.batch_scratch_size(len(entries)), dtype=torch.uint8, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            T.encode_batch_device(batch, t_batch.data_ptr(), st)
        torch.cuda.synchronize()
        ms = []
        for _ in range(args.steps):
            flush.fill_(1)  # L2 flush between timed iterations
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            T.encode_batch_device(batch, t_batch.data_ptr(), st)
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        dev_ms = float(np.mean(ms))
        # ---- end to end: pinned host buffers through the work queue ----
        hin, hout, lens = [], [], []
        for img, d, _ in texs:
            p = L.rosy_tex_alloc_host(img.nbytes)
            C.memmove(p, img.ctypes.data, img.nbytes)
            hin.append(p)
            hout.append(L.rosy_tex_alloc_host(T.dds_size(d)))
            lens.append(C.c_size_t())

        def e2e_step():
            for (img, d, _), pi, po, ln in zip(texs, hin, hout, lens):
                rc = L.rosy_tex_submit_rgba8(C.byref(d), pi, po, T.dds_size(d), C.byref(ln))
                assert rc == 0, L.rosy_tex_last_error()
            assert L.rosy_tex_wait_all() == 0, L.rosy_tex_last_error()

        e2e_step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        # device and e2e legs must agree byte for byte
        d0 = texs[0][1]
        got = np.ctypeslib.as_array(C.cast(hout[0], C.POINTER(C.c_uint8)), shape=(T.dds_size(d0),))
        nb = T.dds_size(d0) - T.DDS_HEADER_BYTES
        same = bool((got[T.DDS_HEADER_BYTES:] == keep[0][3][:nb].cpu().numpy()).all())
        for p in hin + hout:
            L.rosy_tex_free_host(p)
        line = {"config": name, "textures": len(texs), "mpix_level0": mpix, "device_ms": dev_ms,
                "device_mpix_s": mpix / dev_ms * 1e3, "e2e_ms": e2e_ms, "e2e_mpix_s": mpix / e2e_ms * 1e3,
                "e2e_equals_device": same}
        if args.psnr:
            # PSNR delta: per level and chain-weighted, on the FULL image of one texture per kind and size
            # (the oracle is the checker; it is never on the product path).  Each encoder is scored against the
            # uncompressed level it was given (GPU and oracle chains differ by <= 1 LSB).
            import oracle as O

            ps = {}
            for img, d, k in texs:
                key = f"{k}-{img.shape[0]}"
                if key in ps:
                    continue
                h, w = img.shape[:2]
                okind = O.KIND_NORMAL if k == "normal" else O.KIND_SRGB
                fmt, q = int(d.format), int(d.quality)
                g = T.encode_rgba8(img, d.kind, d.format, d.quality)
                o = O.generate_texture(img, okind, fmt, q)
                chain_g, chain_o = T.mips_rgba8(img, d.kind), O.mip_chain(img, okind)
                bb = 8 if fmt == T.FORMAT_BC4 else 16
                ch = {T.FORMAT_BC7: 4, T.FORMAT_BC5: 2, T.FORMAT_BC4: 1}[fmt]
                off, per_level, sg, so, npx, differing, nblk = 148, [], 0, 0, 0, 0, 0
                for lg, lo in zip(chain_g, chain_o):
                    lh, lw = lo.shape[:2]
                    n = ((lw + 3) // 4) * ((lh + 3) // 4) * bb
                    dg = lg[..., :ch].astype(np.int64) - O.decode_image(g[off:off + n], lw, lh, fmt)[..., :ch].astype(np.int64)
                    do = lo[..., :ch].astype(np.int64) - O.decode_image(o[off:off + n], lw, lh, fmt)[..., :ch].astype(np.int64)
                    eg, eo = int((dg * dg).sum()), int((do * do).sum())
                    db = lambda e, m: 99.0 if e == 0 else float(10 * np.log10(255.0 ** 2 * m / e))
                    per_level.append([round(db(eg, dg.size), 3), round(db(eo, do.size), 3)])
                    sg, so, npx = sg + eg, so + eo, npx + dg.size
                    differing += int((g[off:off + n].reshape(-1, bb) != o[off:off + n].reshape(-1, bb)).any(axis=1).sum())
                    nblk += n // bb
                    off += n
                cg = 99.0 if sg == 0 else float(10 * np.log10(255.0 ** 2 * npx / sg))
                co = 99.0 if so == 0 else float(10 * np.log10(255.0 ** 2 * npx / so))
                ps[key] = {"chain_gpu_db": round(cg, 3), "chain_oracle_db": round(co, 3), "chain_delta_db": round(cg - co, 4),
                           "per_level_gpu_oracle_db": per_level,
                           "min_level_delta_db_levels_ge_32px": round(min(a - b for (a, b), l in zip(per_level, chain_o) if l.shape[0] * l.shape[1] >= 1024), 4),
                           "blocks_differing_end_to_end": differing, "blocks": nblk,
                           "max_mip_lsb": int(max(np.abs(a.astype(int) - b.astype(int)).max() for a, b in zip(chain_g, chain_o)))}
            line["psnr_full_image"] = ps
        print(json.dumps(line), flush=True)
        del keep, flush
        torch.cuda.empty_cache()
    T.shutdown()


if __name__ == "__main__":
    main()
