"""Full SASS listings of the hot kernels, one file per kernel, under profiles/sass/ (needs only cuobjdump).
Usage: python bench/sass_listings.py [lib] [outdir]"""
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "dryad_b200/_lib/libdryad_kernels.so"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/sass"
WANT = {  # file stem -> substring(s) of the demangled kernel name
    "sh_scatter_tma": ["sh::sh_scatter_tma"], "sh_scatter": ["sh::sh_scatter("], "sh_count": ["sh::sh_count"],
    "sh_publish": ["sh::sh_publish"], "rs_onesweep_u32_pairs": ["rs_onesweep<unsigned int, (bool)1>"],
    "rs_onesweep_u64_pairs": ["rs_onesweep<unsigned long, (bool)1>"], "gather_rows_fixed_25": ["gather_rows_fixed_kernel<(int)25>"],
    "agg_update_kernel": ["ht::agg_update_kernel<(int)1>", "ht::agg_update_kernel<(int)2>"], "agg_merge_peers_kernel": ["agg_merge_peers_kernel"],
    "km_assign_pair_8": ["km_assign_pair<(int)8>"], "km_certify": ["km::km_certify"], "km_recheck": ["km::km_recheck"],
    "km_update": ["km::km_update"], "nvls_allreduce_f32": ["allreduce_f32"], "nvls_broadcast16": ["broadcast16"],
    "join_probe_unique": ["join_probe_unique_kernel"], "tok_write": ["tok_write_kernel"], "fix_ties32_keys": ["fix_ties32_keys_kernel"],
    "scan_i64_block": ["scan_i64_block_kernel"], "blockchan_checksum": ["bc::checksum_kernel"], "ring_push": ["ring::push_kernel"],
}
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks, name, cur = {}, None, []
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name:
            blocks[name] = cur
        name, cur = m.group(1), [line]
    elif name:
        cur.append(line)
if name:
    blocks[name] = cur
names = list(blocks)
pretty = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
os.makedirs(out, exist_ok=True)
index = []
for stem, subs in WANT.items():
    hits = [(m, p) for m, p in zip(names, pretty) if any(s in p for s in subs)]
    if not hits:
        index.append(f"{stem}: NOT FOUND")
        continue
    with open(os.path.join(out, stem + ".sass"), "w") as f:
        for m, p in hits:
            f.write(f"// {p}\n")
            f.write("\n".join(blocks[m]) + "\n\n")
    index.append(f"{stem}.sass: " + "; ".join(p.split("(")[0] for _, p in hits) + f" ({sum(len(blocks[m]) for m, _ in hits)} lines)")
with open(os.path.join(out, "INDEX.txt"), "w") as f:
    f.write(f"Full SASS (cuobjdump -sass, sm_100a) of the hot kernels of {lib}\n" + "\n".join(index) + "\n")
print("\n".join(index))
