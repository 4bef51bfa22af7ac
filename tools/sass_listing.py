"""Per-kernel SASS listings for profiles/sass/ (cuobjdump -sass of the built library, no GPU needed).

A whole GEMM kernel is several thousand instructions, so a listing keeps what shows how the kernel works: every
tensor-core / TMA / tensor-memory / mbarrier / multimem / reduction / barrier instruction and every branch with its
target, each with its address, plus per-kernel totals - enough to follow the main loops (producer: SYNCS wait ->
UTMALDG; MMA issuer: SYNCS wait -> UTC*MMA x4 -> UTCBAR; epilogue: LDTM -> math -> STS -> UTMASTG).  Small kernels
(collectives, LM-head helpers) are listed in full.

    python tools/sass_listing.py            # writes profiles/sass/<kernel>.sass.txt
"""
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
so = ROOT / "distributed_pytorch_cookbook_b200" / "_C" / "libb200gpt.so"
out_dir = ROOT / "profiles" / "sass"
out_dir.mkdir(parents=True, exist_ok=True)

# (file stem, substring(s) of the demangled name, full listing?)
WANT = [
    ("gemm_pair_plain", ["gemm_tcgen05_kernel<256, 6, 1, 0, 2, false, false>"], False),
    ("gemm_pair_lmhead_softmax_epilogue", ["gemm_tcgen05_kernel<256, 5, 1, 256, 2, false, false>"], False),
    ("gemm_pair_wgrad_reduce_add", ["gemm_tcgen05_kernel<256, 6, 1, 96, 2, false, false>"], False),
    ("gemm_pair_wgrad_grouped", ["gemm_tcgen05_kernel<256, 6, 1, 96, 2, false, true>"], False),
    ("gemm_pair_fused_allgather", ["gemm_tcgen05_kernel<256, 5, 1, 0, 2, true, false>"], False),
    ("gemm_single_fused_allgather", ["gemm_tcgen05_kernel<256, 4, 1, 0, 1, false, false>"], False),
    ("layernorm_bwd_handoff", ["layernorm_bwd_small_kernel<2, 1, true>"], True),
    ("attn_fwd_warp_specialised_d128", ["attn_fwd_ws_kernel<128>"], False),
    ("attn_bwd_warp_specialised_d128", ["attn_bwd_ws_kernel<128, true>"], False),
    ("attn_fwd_per_tile_d32", ["attn_fwd_kernel<32>"], False),
    ("attn_bwd_per_tile_d32", ["attn_bwd_kernel<32>"], False),
    ("allreduce_nvls", ["allreduce_kernel<true>"], True),
    ("allreduce_bf16_wire", ["allreduce_bf16_kernel"], True),
    ("reduce_scatter_nvls", ["reduce_scatter_kernel<true>"], True),
    ("rs_adamw_nvls", ["rs_adamw_kernel<true>"], True),
    ("allgather_pull", ["allgather_kernel"], True),
    ("lm_finish", ["lm_finish_kernel"], True),
    ("adamw", ["adamw_kernel"], True),
]
KEEP = re.compile(r"UTC|UTMA|LDTM|STTM|SYNCS|MULTIMEM|\.MC|REDG|RED\.|ATOM|BAR\.|BRA|EXIT|MUFU|UBLKCP|UBLKRED|CCTL|MEMBAR|FENCE|ELECT|LDGSTS|NANOSLEEP")

text = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
funcs, name, buf = {}, None, []
for line in text.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name is not None:
            funcs[name] = buf
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        buf = []
    elif name is not None:
        buf.append(line)
if name is not None:
    funcs[name] = buf

index = []
for stem, keys, full in WANT:
    hit = next((n for n in funcs if all(k in n for k in keys) and n.replace("void ", "").startswith(keys[0][:12])), None)
    if hit is None:
        index.append(f"{stem}: NOT FOUND ({keys})")
        continue
    ins = [l for l in funcs[hit] if re.search(r"/\*[0-9a-f]{4}\*/", l)]
    code = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in ins]
    counts = {}
    for l in code:
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", l)
        if m:
            op = m.group(1).split(".")[0]
            counts[op] = counts.get(op, 0) + 1
    top = sorted(counts.items(), key=lambda kv: -kv[1])
    head = [f"kernel : {hit.split('(')[0]}", f"source : cuobjdump -sass {so.name} (sm_100a), {len(code)} instructions",
            "opcode totals : " + "  ".join(f"{k}={v}" for k, v in top[:28]), ""]
    body = code if full else [l for l in code if KEEP.search(l)]
    if not full:
        head.insert(2, f"listing: {len(body)} of {len(code)} instructions - tensor-core / TMA / TMEM / mbarrier / multimem / "
                       "reduction / barrier / branch lines, with their addresses")
    (out_dir / f"{stem}.sass.txt").write_text("\n".join(head + body) + "\n")
    index.append(f"{stem}.sass.txt  {hit.split('(')[0]}  ({len(body)} lines)")
(out_dir / "README.md").write_text("# SASS listings (tools/sass_listing.py)\n\n" + "\n".join(f"* {l}" for l in index) + "\n")
print("\n".join(index))
