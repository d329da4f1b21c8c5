"""Per-source-line profile of one kernel: ncu's per-SASS-instruction counters (source page of an `ncu --set full
--import-source on` report) correlated with `nvdisasm -g` of the library the report was taken with.

  python benchmarks/profile_by_line.py gpurun_out/prof.ncu-rep blend_backward 'blend_backward_kernelILi6ELi8ELi1E' \
      activesplat_b200/csrc/blend.cu profiles/rN_blend_backward_by_source_line.txt

argv: report, kernel-name regex for ncu, mangled-name fragment for nvdisasm, source file, output.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, kregex, mangled, src_path, out_path = sys.argv[1:6]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kregex],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = next(i for i, r in enumerate(rows) if "Address" in r)
    h = rows[hdr]
    ix = {n: i for i, n in enumerate(h)}
    inst = {}
    for r in rows[hdr + 1:]:
        if len(r) < len(h):
            continue
        try:
            a = int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]])
            if a not in inst:   # the first launch of the kernel in the report
                inst[a] = (float(r[ix["Instructions Executed"]]), float(r[ix["Thread Instructions Executed"]]),
                           float(r[ix["# Samples"]]), float(r[ix["L1 Wavefronts Shared"]]))
        except ValueError:
            pass
    base = min(inst)
    d = tempfile.mkdtemp()
    lib = os.path.join(ROOT, "activesplat_b200", "libactivesplat_b200.so")
    stem = os.path.basename(src_path).split(".")[0]
    subprocess.run(f"cd {d} && cuobjdump -xelf {stem} {lib} >/dev/null 2>&1 && nvdisasm -g -c {stem}*.cubin > k.sass", shell=True)
    lines = open(os.path.join(d, "k.sass")).read().split("\n")
    start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and mangled in l)
    end = next((i for i, l in enumerate(lines) if l.startswith(".text.") and i > start), len(lines))
    per = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])
    cur = None
    for l in lines[start:end]:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"^\s*/\*([0-9a-f]+)\*/", l)
        if m and (base + int(m.group(1), 16)) in inst:
            v = inst[base + int(m.group(1), 16)]
            for i in range(4):
                per[cur][i] += v[i]
    tot = sum(v[0] for v in per.values())
    ts = sum(v[2] for v in per.values())
    tw = sum(v[3] for v in per.values())
    src = open(src_path).read().split("\n")
    out = [f"# {kregex} ({mangled}): warp instructions per source line of {src_path}",
           f"# ncu --set full --import-source on ({rep}, source page) correlated with nvdisasm -g of the same build; inst = share "
           "of warp instructions, samp = share of stall samples, smem = share of shared-memory wavefronts, thr = active threads per instruction",
           "total warp inst %.1fM samples %d shared wavefronts %.1fM" % (tot / 1e6, ts, tw / 1e6)]
    for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])[:70]:
        own = k and k[0] == os.path.basename(src_path) and k[1] <= len(src)
        text = src[k[1] - 1].strip()[:100] if own else "(inlined from %s)" % (k[0] if k else "?")
        out.append("%6.2f%% inst  %5.1f%% samp  %5.1f%% smem  thr %4.1f  L%-4d %s" % (
            100 * v[0] / tot, 100 * v[2] / max(ts, 1), 100 * v[3] / max(tw, 1), v[1] / max(v[0], 1), k[1] if k else -1, text))
    open(out_path, "w").write("\n".join(out) + "\n")
    print("\n".join(out[:30]))


if __name__ == "__main__":
    main()
