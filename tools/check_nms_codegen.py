"""Build-time check of the rotated-NMS arithmetic (needs nvcc + nvdisasm, no GPU).

nvcc decides per call site which products of an expression are fused into FMAs, and that choice decides the last bits of
an IoU.  k_nms_mask<true> is validated bit for bit against the reference's CUDA kernel on the GPU
(tests/test_nms_gpu.py); this script checks that the other kernels that inline the same per-pair code
(k_nms_greedy<true>: two inlined copies; k_nms_prep: the per-box part) got the SAME mix of FMUL / FFMA / FADD on every
source line of that code, i.e. the same contraction.  Exit code 1 on a difference.

    python tools/check_nms_codegen.py
"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "fcaf3d_b200", "csrc", "nms.cu")


def op_table():
    with tempfile.TemporaryDirectory() as d:
        cubin = os.path.join(d, "nms.cubin")
        subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-cubin",
                        "-o", cubin, SRC], check=True, capture_output=True)
        sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], check=True, capture_output=True, text=True).stdout
    kernels, cur, line = collections.OrderedDict(), None, None
    for l in sass.split("\n"):
        m = re.match(r"\.text\.(\S+):", l)
        if m:
            cur, line = kernels.setdefault(m.group(1), collections.Counter()), None
            continue
        m = re.search(r'//## File ".*", line (\d+)', l)
        if m:
            line = int(m.group(1))
            continue
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?(FFMA|FMUL|FADD)[.\w]*\s", l)
        if m and cur is not None and line:
            cur[(line, m.group(1))] += 1
    return kernels


def main():
    src = open(SRC).read().split("\n")
    find = lambda s: next(i for i, l in enumerate(src) if s in l) + 1
    pair_lo, pair_hi = find("float cross3("), find("float iou_aligned(")          # per-pair arithmetic
    prep_lo, prep_hi = find("void rotate_about("), find("int seg_intersect(")      # per-box arithmetic
    # lines whose attribution may shift between each other: corner_in_box and the corner loops of bev_overlap that inline it
    inbox_lo, inbox_hi = find("int corner_in_box("), find("float bev_overlap(")
    loop_lo, loop_hi = find("for (int k = 0; k < 4; ++k) {"), find("ctr.x /= cnt;")
    k = op_table()
    get = lambda pat: next(v for n, v in k.items() if pat in n)
    mask, prep = get("k_nms_maskILb1E"), get("k_nms_prep")
    greedies = [(n, v) for n, v in k.items() if "k_nms_greedyILb1E" in n]      # every thread-count instantiation
    ops = ("FMUL", "FFMA", "FADD")
    bad = []
    for gname, greedy in greedies:
        tag = "greedy<%s>" % re.search(r"ILb1ELi(\d+)E", gname).group(1)
        inbox = {name: collections.Counter() for name in ("mask", "greedy")}
        for ln in range(pair_lo, pair_hi):
            if prep_lo <= ln < prep_hi:
                continue
            for op in ops:
                m, g = mask.get((ln, op), 0), greedy.get((ln, op), 0)
                if inbox_lo <= ln < inbox_hi or loop_lo <= ln < loop_hi:
                    inbox["mask"][op] += m
                    inbox["greedy"][op] += g
                elif g != 2 * m:
                    bad.append((tag, ln, op, m, g))
        # the corner-in-box lines are compared as a group (the same FMA may be attributed to a neighbouring line), and
        # nvcc may unroll the 4-corner loops of one call site, so the group must hold a whole number (>= 2) of copies of
        # the mask kernel's mix - the same multiple for every opcode, i.e. the same contraction in every copy
        ratios = set()
        for op in ops:
            m, g = inbox["mask"][op], inbox["greedy"][op]
            if m == 0:
                if g:
                    bad.append((tag + " corner_in_box", 0, op, m, g))
            elif g % m or g // m < 2:
                bad.append((tag + " corner_in_box", 0, op, m, g))
            else:
                ratios.add(g // m)
        if len(ratios) > 1:
            bad.append((tag + " corner_in_box mix", 0, "ratio", 0, len(ratios)))
    for ln in range(prep_lo, prep_hi):
        if "hx" in src[ln - 1] or "/ 2" in src[ln - 1]:
            continue  # half extents: x * 0.5 is exact, so fusing it into the following add cannot change a bit
        for op in ops:
            m, p = mask.get((ln, op), 0), prep.get((ln, op), 0)
            if 2 * p != m:   # the mask kernel holds two inlined copies (row box, column box)
                bad.append(("prep", ln, op, m, p))
    for b in bad:
        print("MISMATCH %s line %d %s: mask=%d other=%d | %s" % (b + (src[b[1] - 1].strip()[:70] if b[1] else "",)))
    print("rotated-NMS codegen check: %s (%d source lines compared)" % ("FAILED" if bad else "ok", pair_hi - pair_lo))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
