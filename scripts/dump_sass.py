"""Write one SASS listing per kernel family into profiles/sass/ (runs on the build box, no GPU needed).

    python scripts/dump_sass.py            # after python -m llmss_b200.ops.build

`cuobjdump -sass` of the whole library is ~190 MB (108 kernel instantiations); committed are the bf16
instantiations the flagship decode / prefill steps launch, with the instruction encodings stripped (mnemonics and
operands kept), plus `profiles/sass_census.txt`: for every instantiation the count of the mnemonics that prove
the Blackwell path (UTCHMMA* = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA, UTCBAR = tcgen05.commit,
LDGMC = multimem.ld_reduce, SYNCS = mbarrier) and of the legacy tensor path (HMMA: must be 0).
"""
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llmss_b200", "_C", "libllmss_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")

# representative instantiations: (file stem, regex on the demangled name)
PICK = [
    ("gemm_swapab_bf16_bn64", r"gemm_swapab_kernel<__nv_bfloat16, 64, false>"),
    ("gemm_swapab_bf16_bn256", r"gemm_swapab_kernel<__nv_bfloat16, 256, false>"),
    ("gemm_cluster_bf16_bn64", r"gemm_cluster_kernel<__nv_bfloat16, 64, false>"),
    ("gemm_cluster_allreduce_bf16_bn64", r"gemm_cluster_kernel<__nv_bfloat16, 64, true>"),
    ("gemm_pair_2cta_bf16", r"gemm_pair_kernel<__nv_bfloat16>"),
    ("attn_decode_tc_bf16_d256_r1", r"attn_decode_tc_kernel<__nv_bfloat16, 256, 1>"),
    ("attn_decode_tc_bf16_d128_r4", r"attn_decode_tc_kernel<__nv_bfloat16, 128, 4>"),
    ("attn_decode_cudacore_bf16_d256", r"attn_decode_kernel<__nv_bfloat16, 256, 1, 4>"),
    ("attn_prefill_bf16_d256", r"attn_prefill_kernel<__nv_bfloat16, 256, false>"),
    ("attn_prefill_bf16_d128", r"attn_prefill_kernel<__nv_bfloat16, 128, false>"),
    ("attn_prefill_paged_bf16_d128", r"attn_prefill_kernel<__nv_bfloat16, 128, true>"),
    ("norm_residual_bf16", r"norm_residual_kernel<__nv_bfloat16>"),
    ("rope_kv_write_bf16", r"rope_kv_write_kernel<__nv_bfloat16>"),
    ("embed_bf16", r"embed_kernel<__nv_bfloat16>"),
    ("sample", r"::sample_kernel"),
    ("vp_summary", r"::vp_summary_kernel"),
    ("vp_merge_sample", r"::vp_merge_sample_kernel"),
    ("vp_multinomial_pick", r"::vp_multinomial_pick_kernel"),
    ("argmax_local", r"::argmax_local_kernel"),
    ("allreduce_oneshot_bf16", r"allreduce_oneshot_kernel<__nv_bfloat16>"),
    ("allreduce_twoshot_lamport_bf16", r"allreduce_twoshot_lam_kernel<__nv_bfloat16>"),
    ("allreduce_twoshot_flags_bf16", r"allreduce_twoshot_kernel<__nv_bfloat16>"),
    ("allreduce_nvls_bf16_lamport", r"allreduce_nvls_kernel<__nv_bfloat16, true>"),
    ("allreduce_nvls_bf16_flags", r"allreduce_nvls_kernel<__nv_bfloat16, false>"),
    ("allgather", r"::allgather_kernel"),
]
CENSUS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "LDGMC", "SYNCS", "HMMA", "LDGSTS", "ELECT"]


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"(?m)^\s*Function : ", txt)[1:]
    names = subprocess.run(["c++filt"], input="\n".join(p.split("\n", 1)[0].strip() for p in parts), capture_output=True,
                           text=True, check=True).stdout.splitlines()
    strip = re.compile(r"\s*/\* 0x[0-9a-f]+ \*/\s*$")
    census_lines = []
    written = set()
    for name, body in zip(names, parts):
        lines = [strip.sub("", ln) for ln in body.split("\n")[1:]]
        lines = [ln for ln in lines if ln.strip()]
        ops = Counter()
        for ln in lines:
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m:
                for key in CENSUS:
                    if m.group(1).startswith(key):
                        ops[key] += 1
        short = re.sub(r"\(.*", "", name.replace("void (anonymous namespace)::", ""))
        census_lines.append(f"{short:70s} " + " ".join(f"{k}={ops[k]}" for k in CENSUS if ops[k]) + f" | {len(lines)} lines")
        for stem, rx in PICK:
            if stem not in written and re.search(rx, name):
                with open(os.path.join(OUT, stem + ".sass"), "w") as f:
                    f.write(f"// {name}\n// cuobjdump -sass libllmss_b200.so (sm_100a), encodings stripped\n")
                    f.write("\n".join(lines) + "\n")
                written.add(stem)
    with open(os.path.join(ROOT, "profiles", "sass_census.txt"), "w") as f:
        f.write("# mnemonic census of every kernel instantiation in libllmss_b200.so (scripts/dump_sass.py)\n")
        f.write("\n".join(sorted(census_lines)) + "\n")
    missing = [s for s, _ in PICK if s not in written]
    print(f"wrote {len(written)} listings to {OUT}; census of {len(parts)} functions" + (f"; MISSING {missing}" if missing else ""))
    return 1 if missing else 0


if __name__ == "__main__":
    sys.exit(main())
