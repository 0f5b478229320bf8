"""Static evidence from the built library (no GPU needed): per-kernel registers / static shared memory /
stack from `cuobjdump --dump-resource-usage`, and the SASS mnemonics that show which memory path each
kernel uses (16-byte LDG, LDGSTS = cp.async, UBLKCP + SYNCS = TMA bulk copy + mbarrier, RED/ATOM, local
memory).  Writes profiles/r2_resources.md.    python scripts/resources.py"""
import collections, os, re, subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dlrm-flexflow_b200", "libffemb.so")


def run(*cmd):
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def family(demangled):
    n = demangled.replace("(anonymous namespace)::", "").replace("ffemb::", "")
    n = re.sub(r"<.*", "", n.split("(")[0]).split(" ")[-1]
    return re.sub(r"cub::CUB_\d+_SM_\d+::", "cub::", n)


def demangle(names):
    return run("c++filt", *names).splitlines()


res = collections.OrderedDict()
usage = run("cuobjdump", "--dump-resource-usage", LIB)
for blk in usage.split("Fatbin elf code:")[1:]:
    if "arch = sm_100a" not in blk:
        continue
    found = re.findall(r"Function (\S+):\n\s+(REG:\d+.*)", blk)
    for (_, r), n in zip(found, demangle([f for f, _ in found])):
        d = dict(kv.split(":") for kv in r.split())
        e = res.setdefault(family(n), dict(n=0, reg=[], smem=[], stack=[]))
        e["n"] += 1
        e["reg"].append(int(d["REG"])); e["smem"].append(int(d["SHARED"])); e["stack"].append(int(d["STACK"]))

PAT = collections.OrderedDict([
    ("LD(G).128", r"\b(LDG|LD)\.E(\.\w+)*\.128"), ("ST(G).128", r"\b(STG|ST)\.E(\.\w+)*\.128"), ("LDGSTS", r"\bLDGSTS"),
    ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("ATOM.OR", r"\bATOM\.E\.OR"), ("REDG.ADD.F32", r"\bREDG\.E\.ADD\.F32"),
    ("LDL/STL", r"\b(LDL|STL)\b"), ("CALL", r"\bCALL\.")])
sass = run("cuobjdump", "-sass", LIB)
blocks = re.split(r"\n\s*Function : ", sass)[1:]
cnt = collections.OrderedDict()
for b, n in zip(blocks, demangle([b.split("\n")[0].strip() for b in blocks])):
    e = cnt.setdefault(family(n), collections.Counter())
    e["instr"] += len(re.findall(r"/\*[0-9a-f]{4}\*/", b))
    for k, p in PAT.items():
        e[k] += len(re.findall(p, b))

out = ["# Round 2 — static resource usage and memory-path mnemonics of every kernel in libffemb.so",
       "",
       "Made by `python scripts/resources.py` from the library as built by `__graft_entry__.build()` (`nvcc -gencode",
       "arch=compute_100a,code=sm_100a -lineinfo`): `cuobjdump --dump-resource-usage` and `cuobjdump -sass`, summed over",
       "the template instantiations of a kernel.  No GPU involved: this is what the compiler produced, not a measurement.",
       "",
       "| kernel | instantiations | registers (min–max) | static smem (max, B) | stack (max, B) | SASS instr | "
       + " | ".join(PAT) + " |", "|---|---|---|---|---|---|" + "---|" * len(PAT)]
for k, e in res.items():
    if not k:
        continue
    c = cnt.get(k, collections.Counter())
    out.append(f"| `{k}` | {e['n']} | {min(e['reg'])}–{max(e['reg'])} | {max(e['smem'])} | {max(e['stack'])} | {c['instr']} | "
               + " | ".join(str(c[p]) for p in PAT) + " |")
out += ["",
        "Reading it:",
        "* `fwd_kernel` / `reduce1_ring_kernel` / `reduce_*` / `combine_kernel`: rows move as 16-byte `LDG.E…128` / `STG.E…128`",
        "  (`LD.E.128` / `ST.E.128` where the access carries an L2 cache-policy operand); level 1's output-grad ring is `LDGSTS` (`cp.async`); the bag-major small-table kernels stage",
        "  their gradient tiles with `LDGSTS` too.",
        "* `push_bulk_kernel`, `hot_sum_kernel`, `fwd_bulk_kernel`: `UBLKCP` (`cp.async.bulk`, the TMA unit) with",
        "  `SYNCS` (mbarrier) — the push and the hot-row slices are where the unit is a contiguous tile; the bulk",
        "  forward is the measured comparison that lost (`r2_tma_ab.md`).",
        "* `ATOM.OR`: the touched-row bitmap of the full-optimizer modes (`APPLY_SGDX`: one bit per updated row,",
        "  `ffemb_internal.cuh`), compiled into those instantiations only; `REDG.E.ADD.F32x4` (float atomics on",
        "  table rows) exists in `bwd_atomic_kernel` alone — the reference-equivalent comparison.  The product's",
        "  gradient path has no floating-point atomics.",
        "* Stack: the AVG-literal instantiations of `fwd_kernel` call the exact-quotient subroutine (one non-inlined",
        "  call per fragment, DESIGN §3) and keep the fragment on the stack across it; the SUM instantiations of the",
        "  benched shapes (`<32,1,0,4,…>`) have 0–7 local-memory instructions, none in the gather loop.  `reduce1_kernel`",
        "  (stack 88 B) is the non-ring fallback, not on the benched path; `hot_sum_kernel` sits at 224 registers",
        "  (opt-in path, `FFEMB_HOT=1`).",
        ""]
open(os.path.join(ROOT, "profiles", "r2_resources.md"), "w").write("\n".join(out))
print("\n".join(out[6:20]))
