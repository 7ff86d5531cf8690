"""Warp-instruction and stall-sample shares of check_kernel by source region (from an ncu --set full
capture with -lineinfo):  python profiles/regions.py gpurun_out/r01_prof_check_ur5e.ncu-rep"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:check_kernel", "-c", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
sections, cur = [], None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = {"file": r[1], "rows": [], "hdr": None}; sections.append(cur)
    elif r and r[0] == "Line No" and cur is not None:
        cur["hdr"] = r
    elif cur and cur["hdr"] and not (len(r) >= 2 and r[0] == "Function Name"):
        cur["rows"].append(r)
# region boundaries: marker comments inside narrow_phase.cuh; the other files are one region each
src = open("ramp_b200/csrc/narrow_phase.cuh").read().splitlines()
marks = [("narrow: helpers", "// ---- narrow phase over the pair list"), ("stage 1", "// ---- stage 1: bounding spheres"),
         ("stage 1.5", "// ---- stage 1.5"), ("stage 2", "// ---- stage 2"), ("stage 3 (overflow)", "// ---- stage 3"),
         ("kernel body", "__device__ __forceinline__ void reset_results")]
bounds = []
for name, m in marks:
    ln = next(i + 1 for i, t in enumerate(src) if m in t)
    bounds.append((ln, name))
def region(ln):
    name = "narrow: helpers"
    for b, n in bounds:
        if ln >= b: name = n
    return name
tile_src = open("ramp_b200/csrc/tile.cuh").read().splitlines()
fk_line = next(i + 1 for i, t in enumerate(tile_src) if "// Geometry placements are kept as 9 floats" in t)
dm = [(0, "dm: xf / vectors"), (100, "dm: sincos / joints"), (162, "dm: support"), (208, "dm: closed forms / seg-seg"), (262, "dm: obb"), (295, "dm: simplex / gjk")]
inst, samp = collections.Counter(), collections.Counter()
num = lambda x: float(x) if x.replace(".", "", 1).isdigit() else 0.0
for s in sections:
    h = s["hdr"]; iS, iI = h.index("# Samples"), h.index("Instructions Executed")
    fn = s["file"].split("/")[-1]
    for r in s["rows"]:
        if not r[0].isdigit(): continue
        ln = int(r[0])
        if fn == "narrow_phase.cuh": name = region(ln)
        elif fn == "tile.cuh": name = "fk" if ln >= fk_line else "rows / expand_q / tables"
        elif fn == "device_math.cuh": name = [n for b, n in dm if ln >= b][-1]
        else: name = fn
        inst[name] += num(r[iI]); samp[name] += num(r[iS])
I, S = sum(inst.values()), sum(samp.values())
print(f"{'region':32s} warp-inst%  samples%")
for n, v in inst.most_common():
    print(f"{n:32s} {100 * v / I:8.1f} {100 * samp[n] / S:9.1f}")
