"""Annotated SASS of one kernel with the execution counts of an ncu capture: every instruction with its source line, the number
of warp-level executions, the threads active per execution and the stall samples; plus the hottest source lines.

  ncu -i rep.ncu-rep --page source --csv --print-source sass --kernel-name regex:primary_rays > k.csv
  python tools/sass_dyn.py <mangled kernel prefix, e.g. _ZN2vx12primary_raysILi0ELb0ELb0ELb1E> k.csv out.txt

The library must be the one the capture ran (same source, same flags): both list the kernel's instructions in address order
(nvdisasm -g on the cubin inside libvxrt.so gives the line table).  What tools/sass_by_source.py sums per function, this
shows per instruction — it is how the division slow path (zero numerators) and the dead generic-factor code were found."""
import collections, csv, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(prefix):
    tmp = tempfile.mkdtemp(prefix="vxsass")
    lib = os.environ.get("VXRT_LIB", os.path.join(ROOT, "voxelengine_b200", "lib", "libvxrt.so"))
    subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.startswith("vx_trace") and f.endswith(".cubin")][0]
    lines = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], check=True, capture_output=True, text=True).stdout.split("\n")
    start = next(i for i, l in enumerate(lines) if l.startswith(prefix) and l.endswith(":"))
    out, cur = [], ("?", 0)
    for l in lines[start + 1:]:
        if l.startswith("//---------------------"):
            break
        m = re.match(r'\s*//## File "(.*?)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)).replace("vx_", "").split(".")[0], int(m.group(2)))
            continue
        m = re.match(r'\s*/\*([0-9a-f]+)\*/\s+(.*?);', l)
        if m:
            out.append(("%s:%d" % cur, m.group(1), m.group(2).strip()))
        elif l.startswith(".L_"):
            out.append((None, None, l))
    return out


def counts(path):
    rows, hdr = [], None
    for row in csv.reader(open(path)):
        if len(row) > 5 and row[0] == "Address":
            hdr = row
        elif hdr and len(row) == len(hdr) and row[0].startswith("0x"):
            rows.append(dict(zip(hdr, row)))
    return rows


def main():
    prefix, csvf, outf = sys.argv[1], sys.argv[2], sys.argv[3]
    table = line_table(prefix)
    n = sum(1 for t in table if t[0])
    rows = counts(csvf)[:n]  # the CSV lists the kernel twice
    assert len(rows) == n, "the capture has %d instructions, the library %d: not the same build" % (len(rows), n)
    per, thr, total, i, out = collections.Counter(), collections.Counter(), 0, 0, []
    for src, addr, text in table:
        if src is None:
            out.append(text)
            continue
        r = rows[i]; i += 1
        ie, te = int(r["Instructions Executed"]), int(r["Thread Instructions Executed"])
        per[src] += ie; thr[src] += te; total += ie
        out.append("%10d %5.1f %6s  %-16s %s  %s" % (ie, te / ie if ie else 0.0, r["# Samples"], src, addr, text))
    open(outf, "w").write("\n".join(out) + "\n")
    print("%s: %d instructions, %d warp-level executions" % (prefix, n, total))
    for k, v in per.most_common(30):
        print("  %-28s %6.2f %%  %5.1f lanes" % (k, 100.0 * v / total, thr[k] / v))


if __name__ == "__main__":
    main()
