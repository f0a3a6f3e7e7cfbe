#!/usr/bin/env bash
# Full SASS listings (no GPU needed) of the fused paths: one file per kernel under profiles/sass/.
#   scripts/sass_listings.sh
cd "$(dirname "$0")/.."
lib=distributed-machine-learning-project_b200/lib/libdmlp_knn.so
mkdir -p profiles/sass
cuobjdump -sass "$lib" > /tmp/dmlp_all.sass
python3 - <<'PY'
import re
text = open("/tmp/dmlp_all.sass").read()
parts = re.split(r"(?m)^\s*Function : ", text)
want = {"dist_topk_pair_kernelILi1": "dist_topk_pair_kernel_wide", "dist_topk_pair_kernelILi0": "dist_topk_pair_kernel_narrow",
        "dist_topk_pair_kernelILi2": "dist_topk_pair_kernel_skew", "dist_topk_pair_kernelILi3": "dist_topk_pair_kernel_two_epilogues",
        "unpack_bits_kernelILb0": "unpack_bits_kernel_i32", "unpack_bits_kernelILb1": "unpack_bits_kernel_f64", "rerank_emit_kernel": "rerank_emit_kernel",
        "rerank_exact_dedup_kernelILi0": "rerank_exact_dedup_kernel_fp64_rows", "rerank_exact_dedup_kernelILi1": "rerank_exact_dedup_kernel_i32_rows_8B",
        "rerank_exact_dedup_kernelILi2": "rerank_exact_dedup_kernel", "rerank_select_kernel": "rerank_select_kernel",
        "finalize_kernel": "finalize_kernel", "pack_kernelILb0": "pack_kernel", "pack_kernelILb1": "pack_kernel_decimal6", "exact_scan_kernel": "exact_scan_kernel",
        "exact_merge_kernel": "exact_merge_kernel"}
for p in parts[1:]:
    name = p.split("\n", 1)[0].strip()
    for key, out in want.items():
        if key in name:
            body = "\n".join(ln.split("/*", 2)[1].split("*/")[1].rstrip() if ln.count("/*") >= 2 and re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln) else None or ""
                             for ln in p.split("\n")[1:])
            lines = [ln for ln in body.split("\n") if ln.strip()]
            with open(f"profiles/sass/{out}.sass", "w") as f:
                f.write(f"// cuobjdump -sass of {name} (nvcc 12.9, sm_100a); encodings stripped\n")
                f.write("\n".join(lines) + "\n")
            print(out, len(lines), "instructions")
PY
