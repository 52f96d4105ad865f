"""Timeline of the backward kernel's roles from a -DSKP_BWD_TRACE build (csrc/attn_tc_bwd.cu: BW_TR).

  nvcc ... -DSKP_BWD_TRACE -o build/libskp_trace.so stablekeypoints_b200/csrc/*.cu -lcuda
  (GPU box) cp build/libskp_trace.so stablekeypoints_b200/libskp_b200.so; python scripts/bench_kernels.py | grep '^TR' | tail -120 > trace.txt
  python scripts/bwd_trace.py trace.txt

Lane 0 of every warp of CTA 2 stamps clock64 at the marked events of 12 consecutive units; cycles are relative to the S issuer's first stamp.
"""
import collections
import sys

rows = collections.defaultdict(dict)
for line in open(sys.argv[1]):
    p = line.split()
    rows[int(p[1][1:])][int(p[2][1:])] = list(map(int, p[3:]))
fmt = lambda x: f"{x:7d}" if x > -10 ** 9 else "      -"
print("S issuer (warp 1):    start  S_sel(u+1) issued  s_free(u) seen  S(u+1) issued | period")
prev = None
for u in sorted(rows[1]):
    e = rows[1][u]
    print(f"u{u:02d}", *[fmt(x) for x in e[:4]], "|", e[0] - prev if prev is not None else "")
    prev = e[0]
for w in (4, 9):
    print(f"softmax warp {w}:   top  prologue<  prologue>  -  S full  S in regs (s_free)  chunk 0 + prep  held chunks stored  "
          "chunks done  ds_full arrive  mma_done seen  flush staged | period")
    prev = None
    for u in sorted(rows[w]):
        e = rows[w][u]
        print(f"u{u:02d}", *[fmt(e[i]) for i in (0, 1, 2, 3, 4, 7, 5, 6, 8, 9, 11, 10)], "|", e[0] - prev if prev is not None else "")
        prev = e[0]
