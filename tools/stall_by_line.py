"""Warp-stall samples of an `ncu --page source --csv` export, summed per CUDA source line.

usage: python tools/stall_by_line.py <source.csv> <cubin-or-so> <mangled kernel substring> [top]
The SASS of the kernel is disassembled with `nvdisasm -g` (line info from -lineinfo) and joined with the csv by instruction offset."""
import csv
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(binary, kernel):
    tmp = tempfile.mkdtemp()
    cubins = [binary]
    if not binary.endswith(".cubin"):
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(binary)], cwd=tmp, capture_output=True)
        cubins = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")]
    for cb in cubins:
        txt = subprocess.run(["nvdisasm", "-g", "-c", cb], capture_output=True, text=True).stdout
        m = re.search(r"^\.text\.(\S*%s\S*):$" % re.escape(kernel), txt, re.M)
        if not m:
            continue
        body = txt[m.end():]
        nxt = re.search(r"^\.text\.", body, re.M)
        body = body[:nxt.start()] if nxt else body
        cur, out = None, {}
        for l in body.splitlines():
            s = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if s:
                cur = (os.path.basename(s.group(1)), int(s.group(2)))
                continue
            i = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if i:
                out[int(i.group(1), 16)] = (cur, i.group(2))
        return out
    raise SystemExit("kernel not found")


def main():
    src, binary, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    rows = list(csv.reader(open(src)))
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    off = sass_lines(binary, kernel)
    base = int(data[0][ix["Address"]], 16)
    agg = {}
    for r in data:
        o = int(r[ix["Address"]], 16) - base
        line = off.get(o, (("?", 0), ""))[0]
        a = agg.setdefault(line, {"n": 0, "inst": 0, "st": {}})
        a["n"] += int(r[ix["# Samples"]] or 0)
        a["inst"] += int(r[ix["Instructions Executed"]] or 0)
        for k in stalls:
            v = int(r[ix[k]] or 0)
            if v:
                a["st"][k] = a["st"].get(k, 0) + v
    tot = sum(a["n"] for a in agg.values())
    inst = sum(a["inst"] for a in agg.values())
    print(f"samples {tot}, warp instructions {inst}")
    allst = {}
    for a in agg.values():
        for k, v in a["st"].items():
            allst[k] = allst.get(k, 0) + v
    print("stalls:", ", ".join(f"{k[6:]} {v / tot:.3f}" for k, v in sorted(allst.items(), key=lambda x: -x[1])[:8]))
    for line, a in sorted(agg.items(), key=lambda x: -x[1]["n"])[:top]:
        b = sorted(a["st"].items(), key=lambda x: -x[1])[:3]
        print(f"{line[0]}:{line[1]:<5d} {a['n'] / tot * 100:5.1f} %  inst {a['inst'] / inst * 100:5.1f} %  " + ", ".join(f"{k[6:]} {v}" for k, v in b))


if __name__ == "__main__":
    main()
