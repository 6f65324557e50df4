"""Where the warp instructions of one `ncu --set full --import-source on` capture go, by source line and by function:
python scripts/ncu_source_lines.py gpurun_out/v4/prof_r1_final6.ncu-rep [top_n]   (markdown on stdout)

Uses the cuda,sass source page: one aggregated row per source line (instructions executed, thread instructions,
stall samples).  Lines of inlined helpers are attributed to the file / function they were written in."""
import csv, io, os, re, subprocess, sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, cur, hdr = [], None, None
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = os.path.basename(r[1]); continue
    if r[0] == "Line No":
        hdr = r; continue
    if r[0] == "Function Name" or hdr is None or len(r) < len(hdr) or r[2] != "-":
        continue                                             # per-SASS rows carry an address; keep the per-line aggregates
    ie, it, sa = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    try:
        rows.append((cur, int(r[0]), r[1].strip(), int(r[ie]), int(r[it]), int(r[sa])))
    except ValueError:
        pass
tot_i = sum(r[3] for r in rows); tot_t = sum(r[4] for r in rows); tot_s = sum(r[5] for r in rows)

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ray_tracer_b200", "csrc")
def functions(fn):
    """[(first line, name)] of the functions / sections of a source file."""
    out = []
    for n, l in enumerate(open(os.path.join(ROOT, fn)), 1):
        m = re.match(r"\s*(?:static\s+)?(?:__host__\s+)?__device__.*?\b(\w+)\s*\(", l)
        if m and not l.strip().startswith("//"):
            out.append((n, m.group(1)))
        for mark, name in (("SHADE section", "SHADE section (item queue, camera ray, traversal set-up)"), ("node loop: descend", "TRAVERSE: node loop"),
                           ("leaf loop: test parked", "TRAVERSE: leaf loop"), ("lanes with nothing left", "TRAVERSE: exit votes"),
                           ("counters: one atomic", "counters")):
            if mark in l and l.strip().startswith("//") and fn == "rt_trace.cu" and n > 60:
                out.append((n, name))
        if fn == "rt_trace.cu" and l.startswith("trace_kernel("):
            out.append((n, "trace_kernel: set-up, per-lane item state in shared memory"))
    return sorted(out)
cache = {}
def where(fn, line):
    if fn not in cache:
        try: cache[fn] = functions(fn)
        except OSError: cache[fn] = []
    name = "(top of file)"
    for n, f in cache[fn]:
        if n <= line: name = f
    return name

print("total: %.3e warp instructions, %.1f threads per instruction, %d stall samples\n" % (tot_i, tot_t / max(tot_i, 1), tot_s))
agg = {}
for fn, ln, src, i, t, s in rows:
    k = (fn, where(fn, ln)); a = agg.setdefault(k, [0, 0, 0]); a[0] += i; a[1] += t; a[2] += s
print("| file | function / section | warp instructions | share | threads per instruction | stall samples |\n|---|---|---|---|---|---|")
for (fn, f), (i, t, s) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if i * 1000 < tot_i: continue
    print("| `%s` | %s | %.2e | %.1f %% | %.1f | %.1f %% |" % (fn, f, i, 100.0 * i / tot_i, t / max(i, 1), 100.0 * s / max(tot_s, 1)))
print("\n| file:line | source | share of warp instructions | threads per instruction | stall samples |\n|---|---|---|---|---|")
for fn, ln, src, i, t, s in sorted(rows, key=lambda r: -r[3])[:top_n]:
    print("| `%s:%d` | `%s` | %.1f %% | %.1f | %.1f %% |" % (fn, ln, src[:90].replace("|", "\\|"), 100.0 * i / tot_i, t / max(i, 1), 100.0 * s / max(tot_s, 1)))
