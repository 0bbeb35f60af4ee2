"""Counts the SASS mnemonics that prove the Blackwell-native paths, per kernel of libb200km.so (development aid).
usage: python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections, os, re, subprocess, sys

lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "spark_rapids_ml_nai_b200", "lib", "libb200km.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.splitlines()
keys = ("UTCHMMA", "UTMALDG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "SYNCS", "UCGABAR", "FFMA", "DFMA", "DADD", "REDG", "ATOMG", "RED.")
cur, cnt, names = None, collections.OrderedDict(), {}
for ln in sass:
    m = re.search(r"Function : (\S+)", ln)
    if m:
        nm = m.group(1)
        if nm not in names:
            dm = subprocess.run(["cu++filt", nm], capture_output=True, text=True).stdout.strip()
            names[nm] = re.sub(r"\(.*$", "", dm).replace("b200km::<unnamed>::", "").replace("void ", "")
        cur = names[nm]
        cnt.setdefault(cur, collections.Counter())
        continue
    if cur is not None:
        for k in keys:
            if re.search(r"\b" + re.escape(k), ln):
                cnt[cur][k] += 1
print("# SASS mnemonics per kernel (cuobjdump -sass spark_rapids_ml_nai_b200/lib/libb200km.so, sm_100a)\n")
print("UTCHMMA = tcgen05.mma, UTMALDG = cp.async.bulk.tensor (TMA), LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit,")
print("SYNCS = mbarrier operations, UCGABAR = cluster barrier; the `.2CTA` forms (cta_group::2) are in kmeans_gemm_kernel.\n")
print("| kernel | UTCHMMA | UTMALDG | LDTM | STTM | UTCBAR | SYNCS | UCGABAR | FFMA | DFMA+DADD | global atomics / reds |")
print("|---|---|---|---|---|---|---|---|---|---|---|")
for k, c in cnt.items():
    print(f"| `{k}` | {c['UTCHMMA']} | {c['UTMALDG']} | {c['LDTM']} | {c['STTM']} | {c['UTCBAR']} | {c['SYNCS']} | {c['UCGABAR']} | "
          f"{c['FFMA']} | {c['DFMA'] + c['DADD']} | {c['REDG'] + c['ATOMG'] + c['RED.']} |")
two = sum("UTCHMMA.2CTA" in ln for ln in sass)
tma2 = sum("UTMALDG" in ln and "2CTA" in ln for ln in sass)
mc = sum("UTMALDG" in ln and "MULTICAST" in ln for ln in sass)
hm = sum(bool(re.search(r"\bHMMA", ln)) for ln in sass)
print(f"\n`UTCHMMA.2CTA` instructions: {two}; `UTMALDG` with the 2CTA completion form: {tma2}; TMA loads with MULTICAST: {mc}; "
      f"legacy `HMMA` (mma.sync): {hm}.")
