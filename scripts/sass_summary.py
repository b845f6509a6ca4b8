"""Regenerate profiles/sass/: `cuobjdump -sass` of libddl25_kernels.so, one listing per kernel family (first instantiation) and
SUMMARY.txt with the instruction count and the Blackwell mnemonics of every kernel (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld,
UTMALDG = TMA tensor load, UTCBAR = tcgen05.commit, SYNCS = mbarrier, ACQBULK / PREEXIT = griddepcontrol.wait / launch_dependents; HMMA = legacy mma.sync,
expected: none)."""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "ddl25spring_b200", "lib", "libddl25_kernels.so")
out_dir = os.path.join(root, "profiles", "sass")
os.makedirs(out_dir, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
funcs, cur, name = [], [], None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if name:
            funcs.append((name, cur))
        name, cur = m.group(1), []
    elif name:
        if re.match(r"\s*\.{10,}", line) and any("/*0" in l for l in cur):     # end of this function's code
            funcs.append((name, cur))
            name, cur = None, []
            continue
        cur.append(line)
if name:
    funcs.append((name, cur))
dem = subprocess.run(["cu++filt"] + [f[0] for f in funcs], capture_output=True, text=True).stdout.splitlines()
KEYS = ["UTCHMMA", "LDTM", "UTMALDG", "UTMAPF", "UTCBAR", "SYNCS", "BAR", "HMMA", "LDGSTS", "MULTIMEM", "ACQBULK", "PREEXIT", "GRIDDEP"]
lines_out, seen = [], set()
for (mangled, body), d in zip(funcs, dem):
    short = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", d)
    short = re.sub(r"\((int|bool)\)", "", short)
    short = re.sub(r"\(.*", "", short)
    instr = [l for l in body if re.search(r"/\*[0-9a-f]{4}\*/", l)]
    cnt = collections.Counter()
    for l in instr:
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
        if not m:
            continue
        op = m.group(1)
        for k in KEYS:
            if op.startswith(k):
                cnt[k] += 1
    flags = " ".join(f"{k}={cnt[k]}" for k in KEYS if cnt[k])
    lines_out.append(f"{len(instr):6d} instr  {flags:60s} {short}")
    fam = re.sub(r"<.*", "", short).replace("void ", "").strip()
    keep = cnt["UTCHMMA"] or cnt["UTMALDG"] or fam in ("symm_reduce_kernel", "symm_barrier_kernel", "nvls_allreduce_kernel",
                                                       "weighted_sum_kernel", "bn_stats_kernel", "bn_apply_kernel", "krum_select_kernel",
                                                       "rank_select_mean_kernel", "causal_softmax_kernel", "attn_ds_kernel",
                                                       "attn_bwd_prep_kernel", "softmax_xent_wide_kernel", "embedding_bwd_kernel")
    if fam not in seen and keep:        # full listings for the tensor-core / TMA / cross-GPU kernels; the rest is in SUMMARY.txt
        seen.add(fam)
        with open(os.path.join(out_dir, fam + ".sass"), "w") as f:
            f.write(f"// {d}\n" + "\n".join(body) + "\n")
with open(os.path.join(out_dir, "SUMMARY.txt"), "w") as f:
    f.write(__doc__.strip() + "\n\n" + "\n".join(lines_out) + "\n")
print(len(funcs), "kernels ->", out_dir)
