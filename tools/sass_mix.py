"""Static SASS instruction mix of the hot kernels of libyolohead.so (cuobjdump -sass), as markdown.

    python tools/sass_mix.py > profiles/r02_sass_mix.md

Readable without a GPU: how much of each kernel is fp64 (DADD / DMUL / DFMA / DSETP), shared-memory traffic
(LDS / STS / ATOMS), warp collectives (SHFL / VOTE / REDUX / MATCH), and which opcodes dominate.  No tensor-core
(UTC*MMA / HMMA), TMEM (LDTM) or TMA (UTMALDG) opcode is expected in the per-image kernels: nothing there is a
contraction and an image start is only 8-byte aligned.  They are what the fused head Linear + loss kernel (the last
entry, the one contraction next to this path) is made of: UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tensor
load, UTCBAR = tcgen05.commit, SYNCS = mbarrier operations."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "yolo_v1_b200", "libyolohead.so")
WANT = [("N2yh17map_update_kernel", "yh::map_update_kernel"), ("decode_nms_kernelILi1E", "yh::decode_nms_kernel<YH_IOU_INFER>"),
        ("loss_kernelIfLb0E", "yh::loss_kernel<float, staged>"), ("radix_pass", "yh::radix_pass"),
        ("xi_scatter", "yh::xi_scatter"), ("emit_precisions", "yh::emit_precisions"),
        ("head_linear_loss_kernelI13__nv_bfloat16S2_Li2E", "yh::hl::head_linear_loss_kernel<bf16, bf16, 2> (SM pair)")]
GROUPS = {
    "fp64 (DADD DMUL DFMA DSETP)": r"^(DADD|DMUL|DFMA|DSETP|DMNMX)",
    "fp32 (FADD FMUL FFMA FSETP FMNMX FSEL MUFU)": r"^(FADD|FMUL|FFMA|FSETP|FMNMX|FMNMX3|FSEL|MUFU|FCHK)",
    "integer / logic (IMAD IADD3 LOP3 ISETP SEL SHF LEA PRMT VIADD POPC ...)":
        r"^(IMAD|IADD3|IADD|LOP3|ISETP|SEL|SHF|LEA|PRMT|VIADD|POPC|FLO|BREV|IABS|I2F|F2I|I2FP|F2FP|PLOP3|P2R|R2P|VIMNMX|IMNMX|BMSK|SGXT|IDP|VABSDIFF|ULOP3|UMOV|USHF|USEL)",
    "shared memory (LDS STS ATOMS LDSM)": r"^(LDS|STS|ATOMS|LDSM)$",
    "global / local / constant memory (LDG STG ATOMG RED LDL STL LDC)": r"^(LDG|STG|LD|ST|ATOMG|ATOM|RED|LDL|STL|CCTL|LDC|LDCU)$",
    "warp collectives (SHFL VOTE REDUX MATCH WARPSYNC)": r"^(SHFL|VOTE|REDUX|MATCH|WARPSYNC|ENDCOLLECTIVE)",
    "control (BRA BSSY BSYNC CALL RET EXIT BAR NANOSLEEP ...)": r"^(BRA|BSSY|BSYNC|CALL|RET|EXIT|BAR|NANOSLEEP|YIELD|BREAK|JMP|BRX|ERRBAR|MEMBAR|DEPBAR|NOP)",
    "tensor core / TMEM / TMA (HMMA UTC*MMA LDTM UTMALDG ...)": r"^(HMMA|IMMA|DMMA|UTC|LDTM|STTM|UTMA|UBLK|SYNCS)",
}
dump = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", dump)[1:]
print("# Static SASS mix of the hot kernels (round 2; `python tools/sass_mix.py`)\n")
print("Counts are static instructions in the kernel's code, not executed ones (the executed counts are in the ncu\n"
      "summaries next to this file).  Predicates (`@P0`) are stripped; the opcode is the first dotted token.\n")
for needle, title in WANT:
    body = next((f for f in funcs if needle in f.split("\n", 1)[0]), None)
    if body is None:
        print(f"## {title}: not found\n")
        continue
    ops = []
    for line in body.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            ops.append((m.group(1), line.split("*/", 1)[1].strip().rstrip(";")))
    cnt = collections.Counter(o for o, _ in ops)
    total = sum(cnt.values())
    print(f"## {title}  ({total} instructions, mangled `{body.splitlines()[0].strip()[:70]}`)\n")
    print("| group | instructions | share |\n|---|---|---|")
    seen = 0
    left = dict(cnt)
    for name, pat in GROUPS.items():      # an opcode counts for the first group that claims it
        mine = [k for k in left if re.match(pat, k)]
        c = sum(left.pop(k) for k in mine)
        seen += c
        print(f"| {name} | {c} | {100 * c / total:.1f} % |")
    print(f"| other ({' '.join(sorted(left)[:8])}) | {total - seen} | {100 * (total - seen) / total:.1f} % |\n")
    print("top opcodes: " + ", ".join(f"{k} {v}" for k, v in cnt.most_common(18)) + "\n")
    tc = {k: v for k, v in cnt.items() if re.match(GROUPS["tensor core / TMEM / TMA (HMMA UTC*MMA LDTM UTMALDG ...)"], k)}
    if tc:
        print("tensor core / TMEM / TMA opcodes: " + ", ".join(f"{k} {v}" for k, v in sorted(tc.items(), key=lambda kv: -kv[1])) + "\n")
        ex = [t for o, t in ops if o in ("UTCHMMA", "LDTM", "UTMALDG", "UTCBAR")]
        seen_op, keep = set(), []
        for t in ex:
            op = t.split()[0]
            if op not in seen_op:
                seen_op.add(op)
                keep.append(re.sub(r"\s*/\*.*", "", t).strip())
        print("excerpt: `" + "`, `".join(keep) + "`\n")
    f64 = [t for o, t in ops if re.match(r"^(DADD|DMUL|DFMA|DSETP)", o)]
    if f64:
        print("fp64 excerpt (first 6): `" + "`, `".join(re.sub(r"\s*/\*.*", "", t).strip() for t in f64[:6]) + "`\n")
