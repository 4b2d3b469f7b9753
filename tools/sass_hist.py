"""Static SASS instruction histograms of the hottest kernels: cuobjdump -sass lib.so | python tools/sass_hist.py > profiles/rNN_sass_histograms.txt"""
import collections
import re
import subprocess
import sys
want = ["k_sor_half", "k_sor_persistent", "k_ls_sweep_dataflow", "k_p2g_units", "k_advect", "k_g2p", "k_bin_rank_gather", "k_scan_lookback", "k_ex_fill_layers",
        "k_ls_seed", "k_na_divergence"]
cur, hist = None, {}
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        name, cur = m.group(1), None
        if any(w in name for w in want):
            cur = name
            hist[cur] = collections.Counter()
        continue
    if cur:
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            hist[cur][m.group(1).split('.')[0]] += 1
print("# SASS instruction histograms (cuobjdump -sass of libfluidsim3_b200.so, sm_100a, static counts per kernel)")
print("# Blackwell-specific mnemonics to look for (B200_PROFILING.md): UTMALDG/UTMASTG/UBLKCP (TMA), UTC*MMA, LDTM/STTM -- none: every kernel here is a")
print("# stencil / scatter-gather kernel on LDG/STG/LDS/STS, LDGSTS (cp.async) in the level-set tiles; no contraction, so no tensor-core path.")
for k, h in hist.items():
    dem = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip() or k
    print("\n%s\n  total %d: %s" % (dem[:170], sum(h.values()), ", ".join("%s %d" % ab for ab in h.most_common(18))))
