"""SASS evidence (runs here on the CPU box: cuobjdump reads the cross-compiled library).
Writes profiles/sass/<kernel>.sass for the hot kernels (hex encodings stripped) and
profiles/r01_sass_summary.txt = per-kernel counts of the mnemonics that prove the design:
UTCHMMA (tcgen05.mma), UTMALDG (TMA tensor load), UBLKCP (bulk copy), LDTM (tcgen05.ld), HMMA/LDSM
(mma.sync / ldmatrix), SYNCS (mbarrier), UCGABAR / cluster ops, ACQBULK (PDL wait)."""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "chatterbox_vllm_b200", "libcbx_t3.so")
OUT = os.path.join(ROOT, "profiles", "sass")
HOT = ["paged_attn_decode_bal_kernel", "attn_plan_kernel", "paged_attn_decode_kernel", "paged_attn_small_kernel", "paged_attn_prefill_kernel", "gemm_tc_kernel", "gemv_fused_kernel",
       "cfg_sample_kernel", "rmsnorm_kernel", "decode_mega_kernel", "postprocess_tokens_kernel", "linear_f32_kernel"]
KEEP = {"gemm_tc_kernel_int_128_int_1_int_1", "gemm_tc_kernel_int_128_int_2_int_1", "gemm_tc_kernel_int_64_int_4_int_1",
        "gemv_fused_kernel_int_2_int_1_int_1_int_2_int_12", "gemv_fused_kernel_int_2_int_4_int_0_int_3_int_16",
        "paged_attn_decode_bal_kernel", "paged_attn_small_kernel_int_4", "paged_attn_prefill_kernel", "cfg_sample_kernel",
        "rmsnorm_kernel_int_2"}            # one listing per kernel family; the summary covers every instantiation
KEYS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UBLKCP", "UBLKPF", "LDTM", "HMMA", "LDSM", "SYNCS", "UCGABAR", "ACQBULK", "LDG", "STG",
        "LDS", "STS", "FFMA", "MUFU", "ATOMS", "RED", "BAR.SYNC"]

txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
funcs, cur, name = {}, None, None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = m.group(1)
        cur = funcs.setdefault(name, [])
        continue
    if cur is not None and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        ins = re.sub(r"/\*[0-9a-fx]+\*/", "", line).strip().rstrip(";").strip()
        if ins:
            cur.append(ins)
os.makedirs(OUT, exist_ok=True)
dem = subprocess.run(["cu++filt"] + list(funcs), capture_output=True, text=True).stdout.splitlines()
rows = []
for mangled, d in zip(funcs, dem):
    ins = funcs[mangled]
    short = next((h for h in HOT if h in d), None)
    if short is None:
        continue
    targs = re.search(short + r"<([^>]*)>", d)
    tag = short + ("_" + re.sub(r"[^0-9A-Za-z]+", "_", targs.group(1)).strip("_") if targs else "")
    cnt = collections.Counter()
    for i in ins:
        op = i.split()[1] if i.startswith("@") and len(i.split()) > 1 else i.split()[0]
        for k in KEYS:
            if op.startswith(k):
                cnt[k] += 1
    rows.append((tag, len(ins), cnt))
    if tag not in KEEP:
        continue
    with open(os.path.join(OUT, tag + ".sass"), "w") as f:
        f.write(f"// {d}\n// {len(ins)} instructions, sm_100a, from libcbx_t3.so (cuobjdump -sass)\n")
        f.write("\n".join(ins) + "\n")
with open(os.path.join(ROOT, "profiles", "r01_sass_summary.txt"), "w") as f:
    f.write("SASS mnemonic counts per kernel (tools/dump_sass.py; listings in profiles/sass/)\n")
    f.write(f"{'kernel':58s} {'instr':>6s} " + " ".join(f"{k:>8s}" for k in KEYS) + "\n")
    for tag, n, cnt in sorted(rows):
        f.write(f"{tag:58s} {n:6d} " + " ".join(f"{cnt.get(k, 0):8d}" for k in KEYS) + "\n")
print(open(os.path.join(ROOT, "profiles", "r01_sass_summary.txt")).read())
