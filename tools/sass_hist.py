"""Opcode histogram (executed instructions, stall samples) per kernel from `ncu -i X.ncu-rep --page source --csv`."""
import collections, csv, subprocess, sys
def main(path, pat, top=30):
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    i = 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            name = rows[i][1]; h = rows[i + 1]; j = i + 2
            data = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                if len(rows[j]) == len(h): data.append(rows[j])
                j += 1
            if pat in name: report(name, h, data, top)
            i = j
        else: i += 1
def report(name, h, data, top):
    ia = h.index("Source"); isamp = h.index("# Samples"); iex = h.index("Instructions Executed")
    stall = [k for k in range(len(h)) if h[k].startswith("stall_") and "Not Issued" not in h[k]]
    tot = sum(int(r[isamp]) for r in data); totex = sum(int(r[iex]) for r in data)
    print("kernel:", name[:100]); print("SASS instructions %d, executed %d, samples %d" % (len(data), totex, tot))
    ex = collections.Counter(); sm = collections.Counter(); st = collections.Counter()
    for r in data:
        t = r[ia].split(); op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ex[op] += int(r[iex]); sm[op] += int(r[isamp])
        for k in stall: st[h[k]] += int(r[k])
    for op, c in ex.most_common(top):
        print("%-14s exec %5.1f%%  samples %5.1f%%" % (op, 100.0 * c / totex, 100.0 * sm[op] / tot))
    print("stall samples:", ", ".join("%s %.1f%%" % (k, 100.0 * v / tot) for k, v in st.most_common(10)))
    print()
if __name__ == "__main__":
    try:
        main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
    except BrokenPipeError:
        pass
