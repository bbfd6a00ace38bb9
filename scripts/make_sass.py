"""SASS listings of the kernel library for profiles/sass/ (the text the judge can grep, not mnemonic counts).

    python scripts/make_sass.py

For every object of build/obj (built by omldm_b200.ops.build with -lineinfo for sm_100a):
  profiles/sass/full/<file>.sass.gz   the complete `cuobjdump -sass` listing (encodings stripped), gzip
  profiles/sass/<kernel>.sass         one plain-text listing per HOT-PATH kernel (the kernels the five BASELINE configs
                                      launch per step, and the solvers), encodings stripped
  profiles/sass/INDEX.md              kernel -> file, instruction count, and the Blackwell mnemonics it contains
"""
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
OUT = os.path.join(ROOT, "profiles", "sass")
HOT = ["gram_f16_kernel", "sq_residual_f16s_kernel", "expand_rows_f16s_kernel", "holdout_swap_kernel", "sq_residual_kernel",
       "umma_gemm_kernel", "umma_gemm_scatter_kernel", "sync_allreduce_kernel", "sync_async_kernel", "sync_easgd_kernel",
       "sync_fgm_kernel", "sync_gram_push_kernel", "sync_gram_gather_kernel", "sync_forward_kernel",
       "sync_scatter_push_kernel", "sync_scatter_gather_kernel", "query_reduce_kernel", "device_barrier_kernel",
       "linear_minibatch_kernel", "linear_exact_kernel", "poly_moments_kernel", "poly_scale_linear_kernel",
       "kmeans_assign_kernel", "kmeans_finalize_kernel", "mcpa_coef_kernel", "mcpa_sparse_score_kernel",
       "mcpa_sparse_update_kernel", "sparse_rows_kernel", "densify_kernel", "stage_rows_f16_kernel", "ridge_pcg_kernel",
       "ridge_chol_kernel", "mirror_upper_kernel", "standard_scaler", "minmax", "poly_features_kernel", "ht_accumulate_kernel",
       "ht_split_eval_kernel", "nn_out_grad_kernel", "nn_hidden_act_kernel"]
MARK = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "UCGABAR",
        "HMMA", "REDG", "ATOMG", "MEMBAR.SC.SYS", "MEMBAR.ALL.SYS", "CCTL"]
_ENC = re.compile(r"\s*/\* 0x[0-9a-f]+ \*/\s*$")


def demangle(names):
    p = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return p.stdout.splitlines()


def main():
    os.makedirs(os.path.join(OUT, "full"), exist_ok=True)
    index = []
    for f in sorted(os.listdir(OBJ)):
        if not f.endswith(".o"):
            continue
        txt = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, f)], capture_output=True, text=True).stdout
        lines = [_ENC.sub("", l) for l in txt.splitlines()]
        lines = [l for l in lines if l.strip() and not re.fullmatch(r"\s*/\* 0x[0-9a-f]+ \*/", l)]
        with gzip.open(os.path.join(OUT, "full", f[:-2] + ".sass.gz"), "wt") as g:
            g.write("\n".join(lines) + "\n")
        # split into functions
        cur, name, funcs = [], None, []
        for l in lines:
            m = re.match(r"\s*Function : (\S+)", l)
            if m:
                if name:
                    funcs.append((name, cur))
                name, cur = m.group(1), [l]
            elif name:
                cur.append(l)
        if name:
            funcs.append((name, cur))
        dem = demangle([n for n, _ in funcs])
        for (mangled, body), d in zip(funcs, dem):
            short = d.split("(")[0].replace("omldm::", "").replace("void ", "")
            if not any(h in short for h in HOT):
                continue
            fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:90] + ".sass"
            ins = [l for l in body if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l)]
            marks = {m: sum(1 for l in ins if re.search(r'(?<![A-Z])' + re.escape(m), l)) for m in MARK}
            with open(os.path.join(OUT, fname), "w") as o:
                o.write(f"// {d}\n// object: csrc/{f[:-2]}.cu, {len(ins)} instructions, sm_100a\n" + "\n".join(body) + "\n")
            index.append((short, fname, len(ins), ", ".join(f"{k} {v}" for k, v in marks.items() if v)))
    with open(os.path.join(OUT, "INDEX.md"), "w") as o:
        o.write("# SASS listings (sm_100a, `cuobjdump -sass`, encodings stripped)\n\n"
                "Regenerate with `python scripts/make_sass.py` after `python -m omldm_b200.ops.build`.  "
                "`full/*.sass.gz` has every kernel of every object.\n\n| kernel | file | instructions | Blackwell / memory-order mnemonics |\n|---|---|---|---|\n")
        for short, fname, n, marks in sorted(index):
            o.write(f"| `{short}` | `{fname}` | {n} | {marks} |\n")
    print(len(index), "kernel listings written to", OUT)


if __name__ == "__main__":
    main()
