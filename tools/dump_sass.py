"""Regenerate profiles/sass/: one gzipped SASS listing per hot kernel of the built libfedb200.so and the mnemonic table.

    python tools/dump_sass.py
"""
import gzip, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "federated-learning-with-mpi_b200", "_lib", "libfedb200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# file stem -> substring of the mangled kernel name (template arguments spelled as nvcc mangles them)
KERNELS = [("tcx_round", "k_round_tcx"), ("tcx_fit", "k_fit_tcx"), ("tc_local_round", "k_local_round_tc"), ("gemm_pair", "k_gemm_tma2ILb0"), ("gemm_pair_opt", "k_gemm_tma2ILb1"),
           ("gemm_pair_persistent", "k_gemm_tma2pILb0"), ("gemm_pair_persistent_opt", "k_gemm_tma2pILb1"), ("gemm_tma", "k_gemm_tmaE"),
           ("gemm_cpasync", "k_gemm_tcE"), ("fedavg_p2p", "k_fedavg_p2p"), ("simt_fwd_bwd", "k_fwd_bwd"), ("simt_reduce_opt", "k_reduce_opt"),
           ("simt_fit_cluster", "k_fit_cluster"), ("simt_eval", "k_evalE"), ("simt_fedavg", "k_fedavgE"), ("wide_head_bwd", "k_wide_head_bwd"),
           ("tc_probe", "k_tc_probe")]
COLS = [("UTCHMMA", r"\bUTCHMMA"), ("2CTA", r"\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"), ("UBLKCP", r"\bUBLKCP"),
        ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"), ("UCGABAR", r"\bUCGABAR"), ("LDGSTS", r"\bLDGSTS"), ("*.SYS", r"\.SYS\b")]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
funcs, cur = {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
    elif cur:
        funcs[cur].append(line)
rows = []
for stem, key in KERNELS:
    names = [n for n in funcs if key in n]
    if not names:
        print("missing", key, file=sys.stderr)
        continue
    body = "\n".join(funcs[names[0]])
    with gzip.open(os.path.join(OUT, stem + ".sass.gz"), "wt") as f:
        f.write("Function : " + names[0] + "\n" + body + "\n")
    n_inst = len(re.findall(r"/\*[0-9a-f]{4,5}\*/", body))
    rows.append((stem, names[0], n_inst, [len(re.findall(p, body)) for _, p in COLS]))
with open(os.path.join(OUT, "README.md"), "w") as f:
    f.write("# SASS evidence per hot kernel (cuobjdump -sass of the shipped libfedb200.so, sm_100a; regenerate with tools/dump_sass.py)\n\n"
            "Full listings are the `.sass.gz` files next to this one. Mnemonics that prove the Blackwell-native paths:\n"
            "`UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2: one instruction drives the tensor cores of a CTA pair), `LDTM`/`STTM` = tcgen05.ld/st\n"
            "(TMEM), `UTMALDG` = cp.async.bulk.tensor (TMA), `UBLKCP` = cp.async.bulk, `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier ops,\n"
            "`UCGABAR_*` = cluster barrier, `LDGSTS` = cp.async; `.SYS` = system-scope loads/stores (flags and data in peer GPUs' or host memory).\n\n")
    f.write("| listing | kernel | instructions | " + " | ".join(c for c, _ in COLS) + " |\n|---|---|---|" + "---|" * len(COLS) + "\n")
    for stem, name, n, cnt in rows:
        short = re.sub(r"^_ZN4fedb\d+", "", name)[:40]
        f.write(f"| {stem}.sass.gz | `{short}` | {n} | " + " | ".join(str(c) for c in cnt) + " |\n")
print(open(os.path.join(OUT, "README.md")).read())
