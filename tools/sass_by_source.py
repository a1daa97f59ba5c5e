"""Where a kernel's executed instructions come from, by source function.

Joins the line table of the built library (nvdisasm -g on the cubin inside libvxrt.so) with the per-SASS-instruction
execution counts of an ncu capture (`ncu -i rep --page source --csv --print-source sass --kernel-name regex:K > k.csv`):
both list the kernel's instructions in address order.  Prints, per device function of vx_march.cuh / vx_device.cuh /
vx_trace.cu, its share of the warp instructions executed and the threads active per instruction — once by the function
the line belongs to, once with the small helpers (arithmetic wrappers, cell_bit, ...) folded into the function that
called them.  The library must be the one the capture ran (same source, same flags).

usage: python tools/sass_by_source.py <kernel substring, e.g. shadow_raysILi0ELb0ELb0> k.csv [first last]
       (first/last: also list that range of instructions with their counts)
"""
import collections, csv, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "voxelengine_b200", "csrc")
HELPERS = {"fmul", "fadd", "fsub", "fdiv", "dot3", "mk", "normalize3", "ray_aabb", "cell_bit", "lane_state", "lane_set_state", "lane_dt",
           "lane_ray", "cell_of", "index_of", "interior", "steps_from_start", "to_brick_units", "dda_begin", "dda_inside", "dda_step",
           "normal_from_code", "tiled_index"}


def functions_of(path):
    """[(first line, name)] of the function definitions in a source file (good enough for this code base's layout)."""
    out = []
    pat = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:__device__|__global__|__host__)[^;(]*?\b([A-Za-z_]\w*)\s*\(")
    pending = None
    for n, line in enumerate(open(path), 1):
        if re.match(r"^\s*template\s*<", line) and "(" not in line:
            pending = n
            continue
        m = pat.match(re.sub(r"__launch_bounds__\([^)]*\)", "", line))
        if m:
            out.append((pending or n, m.group(1)))
        pending = None
    return out


def disassemble():
    tmp = tempfile.mkdtemp(prefix="vxsass")
    lib = os.environ.get("VXRT_LIB", os.path.join(ROOT, "voxelengine_b200", "lib", "libvxrt.so"))
    subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.startswith("vx_trace") and f.endswith(".cubin")][0]
    return subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], check=True, capture_output=True, text=True).stdout


def main():
    want, csvf = sys.argv[1], sys.argv[2]
    lo, hi = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1, 0)
    funcs = {f: functions_of(os.path.join(CSRC, f)) for f in ("vx_march.cuh", "vx_device.cuh", "vx_trace.cu")}

    def owner(src):
        if not src:
            return "?"
        f, l = src
        name = None
        for first, n in funcs.get(f, []):
            if first <= l:
                name = n
        return name or f

    cur, inside, lines = None, False, []
    for ln in disassemble().splitlines():
        if ln.startswith(".text."):
            inside = want in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(.*?);", ln)
        if m:
            lines.append((cur, m.group(1)))
    rows = list(csv.reader(open(csvf)))
    hdr = rows[1]
    data = rows[2:2 + len(lines)]
    assert len(data) == len(lines), "capture and library differ (%d vs %d instructions)" % (len(data), len(lines))
    iI, iT = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    own, ownt, ctx, ctxt = collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter()
    total, caller = 0, "?"
    for k, ((src, ins), r) in enumerate(zip(lines, data)):
        c, t = int(r[iI]), int(r[iT])
        total += c
        o = owner(src)
        own[o] += c; ownt[o] += t
        if o not in HELPERS and o != "?" and not o.endswith(".hpp") and not o.endswith(".h"):
            caller = o
        ctx[caller] += c; ctxt[caller] += t
        if lo <= k <= hi:
            print("%5d %11d %5.1f %-14s:%4d  %s" % (k, c, t / max(c, 1), src[0][:14] if src else "-", src[1] if src else 0, ins))
    print("kernel %s: %d SASS instructions, %d warp instructions executed" % (want, len(lines), total))
    for title, a, b in (("by function", own, ownt), ("helpers folded into their caller", ctx, ctxt)):
        print("-- " + title)
        for k, v in a.most_common():
            if v * 1000 >= total:
                print("%5.1f %%  %4.1f lanes  %s" % (100.0 * v / total, b[k] / max(v, 1), k))


if __name__ == "__main__":
    main()
