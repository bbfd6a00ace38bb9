"""Which kernels differ, instruction for instruction, from a given (hardware-validated) commit?

    python scripts/sass_audit.py f92d6d3 > profiles/sass_audit.txt

Builds every csrc/*.cu of that commit and of the working tree for sm_100a and compares the SASS of each kernel by
name.  Used at the end of round 1: work continued after the GPU budget was spent, and this shows that the kernels
measured on hardware are byte-identical in HEAD except where a change was intended."""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]


def build(src_root, out_dir):
    objs = {}
    procs = []
    for f in sorted(os.listdir(os.path.join(src_root, "csrc"))):
        if f.endswith(".cu"):
            o = os.path.join(out_dir, f[:-3] + ".o")
            procs.append((f, o, subprocess.Popen(["nvcc", *FLAGS, "-I", os.path.join(src_root, "include"), "-I",
                                                  os.path.join(src_root, "csrc"), "-c",
                                                  os.path.join(src_root, "csrc", f), "-o", o],
                                                 stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for f, o, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            raise RuntimeError(f"{f}: {out}")
        objs[f] = o
    return objs


def kernels(obj):
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    out, cur = {}, None
    for ln in txt.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            out[cur] = []
        elif cur:
            out[cur].append(ln)
    return out


def demangle(n):
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]


def main():
    ref = sys.argv[1]
    with tempfile.TemporaryDirectory() as td:
        old_src, old_obj, new_obj = os.path.join(td, "src"), os.path.join(td, "old"), os.path.join(td, "new")
        for d in (old_src, old_obj, new_obj):
            os.makedirs(d)
        subprocess.run(f"git -C {ROOT} archive {ref} csrc include | tar -x -C {old_src}", shell=True, check=True)
        a, b = build(old_src, old_obj), build(ROOT, new_obj)
        head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        print(f"# SASS audit: working tree at {head} against {ref} (sm_100a, nvcc flags of ops/build.py)")
        same = changed = new = 0
        for f in sorted(b):
            kb = kernels(b[f])
            ka = kernels(a[f]) if f in a else {}
            for k in sorted(kb):
                if k not in ka:
                    print(f"NEW        {f:22s} {demangle(k)}")
                    new += 1
                elif ka[k] != kb[k]:
                    print(f"CHANGED    {f:22s} {demangle(k)}")
                    changed += 1
                else:
                    same += 1
            for k in sorted(set(ka) - set(kb)):
                print(f"REMOVED    {f:22s} {demangle(k)}   (signature changed or kernel deleted)")
        print(f"# {same} kernels byte-identical, {changed} changed, {new} new")


if __name__ == "__main__":
    main()
