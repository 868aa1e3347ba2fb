"""SASS digest of libslimg_b200.so: per kernel, the counts of the opcodes that prove which hardware paths it uses.
usage: python tools/sass_digest.py > profiles/<round>_sass_digest.txt   (cuobjdump -sass, no GPU needed)"""
import collections, re, subprocess, sys, os
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "slimg_b200", "lib", "libslimg_b200.so")
OPS = ["UTCIMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMAPF", "UBLKCP", "IMMA", "LDSM", "MOVM", "IDP", "I2IP", "LDGSTS", "SYNCS", "FFMA", "STG", "LDG"]
# one instance per template is enough: the C3 / C1 / C4a instances of the hot kernel, the default of the others
KEEP = [r"remap_\w+_kernel", r"resize_fewtaps_kernel<8, 4>", r"resize_generic_kernel", r"resize_fast_kernel<0>",
        r"resize_vfirst_kernel<2, 4, 4, 5>", r"resize_vfirst_kernel<2, 3, 3, 4>", r"resize_vfirst_kernel<2, 2, 2, 3>", r"resize_tc_kernel<4>",
        r"resize_tile_kernel<1, 1, true>", r"jpeg_frontend_warp_kernel<0, 5, false, true, (true|false)>", r"webp_\w+_kernel", r"avif_planes_kernel"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
print("# SASS digest of slimg_b200/lib/libslimg_b200.so (cuobjdump -sass): instruction counts per kernel of the opcodes that prove")
print("# tcgen05 (UTCIMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit), TMA (UTMALDG, UTMAPF prefetch, UBLKCP bulk copy),")
print("# mma.sync (IMMA), ldmatrix (LDSM), movmatrix (MOVM), dp4a (IDP), cvt.pack.sat (I2IP), cp.async (LDGSTS), mbarrier (SYNCS).")
print("# resize_vfirst_kernel<2,4,4,5> is the C3 instance (DEFAULT hot kernel: IMMA + LDSM + UTMALDG + I2IP, no UTCIMMA), <2,3,3,4> C1, <2,2,2,3> C4a;")
print("# resize_tc_kernel is the selectable tcgen05 pipeline (UTCIMMA + LDTM).\n")
for i, body in enumerate(re.split(r"\n\s*Function : \S+\n", sass)[1:]):
    name = re.sub(r"slimg::\(anonymous namespace\)::|void |\(int\)", "", names[i]).replace("(bool)1", "true").replace("(bool)0", "false")
    if not any(re.search(k, name) for k in KEEP): continue
    ops = collections.Counter(m.group(1).split(".")[0] for m in re.finditer(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", body))
    print(name); print("    total %d  " % sum(ops.values()) + "  ".join(f"{o} {ops[o]}" for o in OPS if ops[o]))
