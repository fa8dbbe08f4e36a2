#!/usr/bin/env python3
"""Derive the BC7 (BPTC) partition and anchor tables by probing an independent decoder.

The public BC7 specification defines 64 two-subset partitions, 64 three-subset
partitions and the per-partition "fix-up" (anchor) texels.  No copy of those tables
exists in this image (and none in /root/reference), so instead of typing them from
memory we *measure* them: hand-packed probe blocks are decoded with Pillow's BCn
decoder (PIL/DdsImagePlugin.py, compiled `bcn` decoder) and the subset of every texel
and the position of the shortened index are read back from the decoded colours.

Outputs (identical data, two consumers that must stay independent of each other):
  oracle/bc7_tables.h            - plain C, used by the CPU oracle
  rosy_b200/csrc/bc7_tables.cuh  - __constant__ tables for the sm_100a encoder

Run:  python tools/derive_bc7_tables.py
"""
import io
import os
import struct
import sys

from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Bits:
    def __init__(self):
        self.v = 0
        self.n = 0

    def put(self, val, nbits):
        assert 0 <= val < (1 << nbits)
        self.v |= val << self.n
        self.n += nbits

    def bytes(self):
        assert self.n == 128, self.n
        return self.v.to_bytes(16, "little")


def dds_bc7(blocks, wblocks):
    """Wrap raw BC7 blocks (row of `wblocks` blocks) in a DDS+DX10 container."""
    w, h = 4 * wblocks, 4 * (len(blocks) // wblocks)
    hdr = struct.pack(
        "<4s7I44x", b"DDS ", 124, 0x1 | 0x2 | 0x4 | 0x1000 | 0x80000, h, w, 16 * wblocks, 0, 1
    )
    hdr += struct.pack("<2I4s5I", 32, 0x4, b"DX10", 0, 0, 0, 0, 0)
    hdr += struct.pack("<5I", 0x1000, 0, 0, 0, 0)
    hdr += struct.pack("<5I", 98, 3, 0, 1, 0)
    assert len(hdr) == 148
    return hdr + b"".join(blocks)


def decode(blocks):
    im = Image.open(io.BytesIO(dds_bc7(blocks, len(blocks))))
    im = im.convert("RGBA")
    px = im.load()
    out = []
    for b in range(len(blocks)):
        out.append([px[4 * b + (i & 3), i >> 2] for i in range(16)])
    return out


def mode3_block(partition, index_bits_all):
    """Mode 3: 2 subsets, 7-bit RGB + per-endpoint p-bit, 2-bit indices."""
    b = Bits()
    b.put(0b1000, 4)  # mode 3 = three zero bits then a one (LSB first)
    b.put(partition, 6)
    if index_bits_all == 0:
        # subset 0 black, subset 1 white, all indices 0
        ep = [0, 0, 127, 127]
        pb = [0, 0, 1, 1]
    else:
        ep = [0, 127, 0, 127]
        pb = [0, 1, 0, 1]
    for _ch in range(3):
        for e in ep:
            b.put(e, 7)
    for p in pb:
        b.put(p, 1)
    b.put((1 << 30) - 1 if index_bits_all else 0, 30)
    return b.bytes()


def mode2_block(partition, index_bits_all):
    """Mode 2: 3 subsets, 5-bit RGB, no p-bits, 2-bit indices."""
    b = Bits()
    b.put(0b100, 3)
    b.put(partition, 6)
    if index_bits_all == 0:
        ep = [0, 0, 15, 15, 31, 31]
    else:
        ep = [0, 31, 0, 31, 0, 31]
    for _ch in range(3):
        for e in ep:
            b.put(e, 5)
    b.put((1 << 29) - 1 if index_bits_all else 0, 29)
    return b.bytes()


def derive():
    # ---- 2 subsets ----
    sub2 = decode([mode3_block(p, 0) for p in range(64)])
    part2 = []
    for p in range(64):
        row = []
        for i in range(16):
            r = sub2[p][i][0]
            assert r in (0, 255), (p, i, r)
            row.append(0 if r == 0 else 1)
        assert row[0] == 0
        part2.append(row)
    anc2 = decode([mode3_block(p, 1) for p in range(64)])
    anchor2 = []
    for p in range(64):
        short = [i for i in range(16) if anc2[p][i][0] != 255]
        assert len(short) == 2 and short[0] == 0, (p, short)
        # index 1 of 2-bit ramp between 0 and 255 -> (43*0 + 21*255 + 32) >> 6 = 84
        assert anc2[p][short[1]][0] == 84
        assert part2[p][short[1]] == 1
        anchor2.append(short[1])
    # ---- 3 subsets ----
    sub3 = decode([mode2_block(p, 0) for p in range(64)])
    val2sub = {0: 0, (15 << 3) | (15 >> 2): 1, 255: 2}
    part3 = []
    for p in range(64):
        row = [val2sub[sub3[p][i][0]] for i in range(16)]
        assert row[0] == 0
        part3.append(row)
    anc3 = decode([mode2_block(p, 1) for p in range(64)])
    anchor3a, anchor3b = [], []
    for p in range(64):
        short = [i for i in range(16) if anc3[p][i][0] != 255]
        assert len(short) == 3 and short[0] == 0, (p, short)
        a = [i for i in short[1:] if part3[p][i] == 1]
        c = [i for i in short[1:] if part3[p][i] == 2]
        assert len(a) == 1 and len(c) == 1, (p, short)
        anchor3a.append(a[0])
        anchor3b.append(c[0])
    return part2, anchor2, part3, anchor3a, anchor3b


def fmt(vals, per_line=16, w=2):
    lines = []
    for i in range(0, len(vals), per_line):
        lines.append("    " + ",".join(f"{v:>{w}}" for v in vals[i : i + per_line]) + ",")
    return "\n".join(lines)


def emit(part2, anchor2, part3, anchor3a, anchor3b):
    mask2 = [sum(b << i for i, b in enumerate(r)) for r in part2]  # bit i = subset of texel i
    # 3 subsets: 2 bits per texel
    mask3 = [sum(s << (2 * i) for i, s in enumerate(r)) for r in part3]
    body = f"""\
/* GENERATED by tools/derive_bc7_tables.py - do not edit.
 * BC7 partition / anchor tables, measured by probing Pillow's independent BCn decoder
 * (see the tool's docstring).  Texel i = x + 4*y.
 *   {{P}}PART2_MASK[p]  bit i       = subset (0/1) of texel i in 2-subset partition p
 *   {{P}}PART3_MASK[p]  bits 2i+1:2i = subset (0/1/2) of texel i in 3-subset partition p
 *   {{P}}ANCHOR2[p]     anchor texel of subset 1 (subset 0's anchor is texel 0)
 *   {{P}}ANCHOR3A/B[p]  anchor texels of subsets 1 and 2
 */
"""
    c_hdr = body.replace("{P}", "BC7_") + f"""\
#ifndef ROSY_ORACLE_BC7_TABLES_H
#define ROSY_ORACLE_BC7_TABLES_H
#include <stdint.h>
static const uint16_t BC7_PART2_MASK[64] = {{
{fmt([f'0x{m:04X}' for m in mask2], 8, 6)}
}};
static const uint32_t BC7_PART3_MASK[64] = {{
{fmt([f'0x{m:08X}u' for m in mask3], 4, 11)}
}};
static const uint8_t BC7_ANCHOR2[64] = {{
{fmt(anchor2)}
}};
static const uint8_t BC7_ANCHOR3A[64] = {{
{fmt(anchor3a)}
}};
static const uint8_t BC7_ANCHOR3B[64] = {{
{fmt(anchor3b)}
}};
#endif
"""
    cu_hdr = body.replace("{P}", "c_ / h_") + f"""\
// h_* are host copies (used to derive the byte-mask tables uploaded at init), c_* live in
// constant memory.
#pragma once
#include <stdint.h>
#define ROSY_BC7_PART2_MASK_INIT {{ \\
{fmt([f'0x{m:04X}' for m in mask2], 8, 6).replace(chr(10), ' ' + chr(92) + chr(10))} }}
#define ROSY_BC7_PART3_MASK_INIT {{ \\
{fmt([f'0x{m:08X}u' for m in mask3], 4, 11).replace(chr(10), ' ' + chr(92) + chr(10))} }}
#define ROSY_BC7_ANCHOR2_INIT {{ \\
{fmt(anchor2).replace(chr(10), ' ' + chr(92) + chr(10))} }}
#define ROSY_BC7_ANCHOR3A_INIT {{ \\
{fmt(anchor3a).replace(chr(10), ' ' + chr(92) + chr(10))} }}
#define ROSY_BC7_ANCHOR3B_INIT {{ \\
{fmt(anchor3b).replace(chr(10), ' ' + chr(92) + chr(10))} }}
static const uint16_t h_PART2_MASK[64] = ROSY_BC7_PART2_MASK_INIT;
static const uint32_t h_PART3_MASK[64] = ROSY_BC7_PART3_MASK_INIT;
static const uint8_t h_ANCHOR2[64] = ROSY_BC7_ANCHOR2_INIT;
static const uint8_t h_ANCHOR3A[64] = ROSY_BC7_ANCHOR3A_INIT;
static const uint8_t h_ANCHOR3B[64] = ROSY_BC7_ANCHOR3B_INIT;
__constant__ uint16_t c_PART2_MASK[64] = ROSY_BC7_PART2_MASK_INIT;
__constant__ uint32_t c_PART3_MASK[64] = ROSY_BC7_PART3_MASK_INIT;
"""
    with open(os.path.join(ROOT, "oracle", "bc7_tables.h"), "w") as f:
        f.write(c_hdr)
    with open(os.path.join(ROOT, "rosy_b200", "csrc", "bc7_tables.cuh"), "w") as f:
        f.write(cu_hdr)


if __name__ == "__main__":
    t = derive()
    emit(*t)
    print("2-subset[0..3]:", [hex(sum(b << i for i, b in enumerate(r))) for r in t[0][:4]])
    print("anchor2:", t[1])
    print("anchor3a:", t[3])
    print("anchor3b:", t[4])
    sys.exit(0)
