"""Write one SASS listing per kernel of the training step / collectives to profiles/sass/ (from the built objects,
no GPU needed) plus an index with the tensor-core / TMA / multimem mnemonics each one contains."""
import os, re, subprocess, sys, collections
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(root, "profiles", "sass")
os.makedirs(out, exist_ok=True)
KEEP = re.compile(r"conv1_fwd_tc_kernel|conv1_hz_kernel|conv1_wgrad_tc_kernel|conv1_wgrad_tf32_kernel|conv1_wgrad_tf32_tma_kernel|conv1_wgrad_nox1|conv5x5_rs_kernel.*Lb0E(f|13__nv_bfloat16)Lb1EE|"
                  r"conv5x5_wgrad_rep_kernelI13|conv5x5_wgrad_tf32_kernel|bn_relu_pool_kernel|pool_bn_bwd_kernelILi32ELi1E|"
                  r"fc_fwd_kernelILi5E|fc_bwd_kernelILi5E|fc_bwd_anybatch|fc_reduce|ce_kernel|bn_finalize|bn_bwd_finalize|"
                  r"reduce_partials|sgd_kernel|resize_pil|allreduce_nvls_f32_excl|allreduce_tile_f32_kernelILb1|allreduce_tile_f32_kernelILb0|"
                  r"allreduce_oneshot_f32x4_kernelILi8|allreduce_push_kernelIffE|allreduce_twoshot_f32|barrier_kernel")
MARK = ("UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "LDGSTS", "ARRIVES.LDGSTSBAR", "SYNCS",
        "LDGMC", "STGMC", "MULTIMEM", "RED.E", "STG.E.ENL2.256", "LDG.E.ENL2.256", "FENCE.VIEW.ASYNC", "ELECT")
index = []
for obj in sorted(os.listdir(os.path.join(root, "build"))):
    if not obj.endswith(".o") or obj.endswith(".host.o"):
        continue
    txt = subprocess.run(["cuobjdump", "-sass", os.path.join(root, "build", obj)], capture_output=True, text=True).stdout
    for block in txt.split("\t\tFunction : ")[1:]:
        name, _, body = block.partition("\n")
        if not KEEP.search(name):
            continue
        dem = subprocess.run(["cu++filt", name.strip()], capture_output=True, text=True).stdout.strip() or name.strip()
        short = re.sub(r"^void |\(anonymous namespace\)::|<unnamed>::", "", dem)
        short = re.sub(r"\((int|bool|unsigned int)\)", "", short)      # template-argument casts
        short = re.sub(r"\(.*", "", short)                              # parameter list
        fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:90] + ".sass"
        lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in body.splitlines() if "/*" in l and not re.match(r"^\s*/\* 0x", l)]
        census = collections.Counter()
        for l in lines:
            for m in MARK:
                if re.search(r"\b" + re.escape(m), l):
                    census[m] += 1
        with open(os.path.join(out, fname), "w") as f:
            f.write(f"// {dem}\n// from build/{obj} (nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo); encodings stripped\n")
            f.write("\n".join(lines) + "\n")
        index.append((fname, len(lines), dict(census)))
with open(os.path.join(out, "INDEX.txt"), "w") as f:
    f.write("SASS listings of the kernels on the default paths (tools/dump_sass.py).  instruction count | marker mnemonics\n")
    for fname, n, c in sorted(index):
        f.write(f"{fname:95s} {n:6d}  " + " ".join(f"{k}x{v}" for k, v in sorted(c.items())) + "\n")
print(f"{len(index)} listings -> {out}")
