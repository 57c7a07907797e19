"""Where the 'no instruction' (instruction fetch) stall samples of a profiled kernel sit: per 256-instruction range of the
SASS, with the executed share — and the list of CALL targets (subroutines) with call counts.
    python tools/ncu_noinst.py report.ncu-rep [bucket]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 256
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
isrc, iex, ismp, ini = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("stall_no_inst")
body = [r for r in rows[2:] if len(r) > ini]
tot_ex = sum(int(r[iex] or 0) for r in body); tot_ni = sum(int(r[ini] or 0) for r in body); tot_s = sum(int(r[ismp] or 0) for r in body)
print("instructions", len(body), "executed", tot_ex, "samples", tot_s, "no_inst samples", tot_ni)
ex = collections.Counter(); ni = collections.Counter(); sm = collections.Counter(); touched = collections.Counter()
for n, r in enumerate(body):
    b = n // bucket
    ex[b] += int(r[iex] or 0); ni[b] += int(r[ini] or 0); sm[b] += int(r[ismp] or 0)
    if int(r[iex] or 0) > 0: touched[b] += 1
print("range        exec%  samples%  no_inst%   executed-instr-in-range")
for b in sorted(ex):
    if ex[b] * 200 > tot_ex or ni[b] * 100 > tot_ni:
        print("#%5d-%5d  %5.1f   %5.1f    %5.1f      %d" % (b * bucket, (b + 1) * bucket - 1, 100.0 * ex[b] / tot_ex, 100.0 * sm[b] / tot_s, 100.0 * ni[b] / max(tot_ni, 1), touched[b]))
hot = sum(1 for r in body if int(r[iex] or 0) > 0)
print("distinct instructions executed at least once:", hot, "(%.0f KB)" % (hot * 16 / 1024.0))
top = sorted(range(len(body)), key=lambda i: -int(body[i][ini] or 0))[:12]
for i in top:
    print("  @%5d no_inst %5s exec %9s  %s" % (i, body[i][ini], body[i][iex], body[i][isrc][:70]))
