#!/bin/bash
# Writes profiles/r2_sass_excerpt_headline.txt from the built library: the instructions that show what the
# headline kernel is made of (TMA tensor loads, bulk copies, mbarrier ops, LDS.128, STG.128, elect, Philox).
cd "$(dirname "$0")/.."
K='_ZN3gsb18stencil_tma_kernelIfLi1ELb1ELi8ELi4ELi3ELb1EEEv14CUtensorMap_stS1_PT_S3_NS_6LayoutENS_6ParamsENS_3BoxEiNS_9XFaceArgsE'
S=/tmp/headline_xf.sass
cuobjdump -sass -fun "$K" grayscott.jl_b200/libgsb200.so 2>/dev/null | sed -E 's#\s*/\* 0x[0-9a-f]+ \*/##' | grep -E "^\s+/\*[0-9a-f]{4}\*/" > $S
{
echo "# SASS excerpt of the headline kernel as shipped in grayscott.jl_b200/libgsb200.so (sm_100a):"
echo "#   stencil_tma_kernel<float, ARITH_NATIVE, noise, TY=8, STAGES=4, 3 CTAs/SM, XF=true>   (XF=false: same without UBLKCP)"
echo "#   cuobjdump -sass -fun '$K' grayscott.jl_b200/libgsb200.so      ($(wc -l < $S) instructions)"
echo "#   ptxas: $(grep -A3 "$K" build/csrc/gs_stencil_f32.ptxas.log | grep -E "Used|spill" | tr -s ' ' | tr '\n' ';')"
echo "# Counts over the whole kernel:"
for op in UTMALDG.3D UBLKCP.S.G UBLKCP.G.S SYNCS.EXCH.64 SYNCS.ARRIVE.TRANS64 SYNCS.PHASECHK.TRANS64.TRYWAIT NANOSLEEP ELECT LDS.128 STG.E.EF.128 SHFL IMAD.WIDE.U32 FFMA FADD STL LDL; do printf "#   %-34s %s\n" $op $(grep -c "$op" $S); done
echo; echo "# --- barrier set-up"; grep -m4 "SYNCS.EXCH" $S
echo "# --- producer: ghost column in (bulk copy onto the tile's mbarrier), tile of u and v (TMA tensor loads)"
grep -m2 "UBLKCP.S.G" $S; grep -m2 -B1 "UTMALDG.3D" $S | head -6
echo "# --- producer: boundary column out (bulk store from the slot's side area), commit + wait for the read"
grep -m2 -A3 "UBLKCP.G.S" $S | head -8
echo "# --- consumer: wait for the plane (try_wait with suspend hint -> NANOSLEEP back-off)"
grep -m1 -A3 "PHASECHK.TRANS64.TRYWAIT P6" $S
echo "# --- consumer: centre vectors, y-halo rows and edge scalars from the ring slot"
grep -m6 "LDS" $S
echo "# --- consumer: x neighbours across lanes"; grep -m4 "SHFL" $S
echo "# --- consumer: Philox round (32x32->64 multiply, 3-input xor with the round key from a uniform register)"
grep -m3 -A1 "IMAD.WIDE.U32" $S | head -6
echo "# --- consumer: streaming 16-byte stores, boundary column into the side area, slot release by one elected lane"
grep -m2 "STG.E.EF.128" $S; grep -m2 "STS" $S; grep -m1 -B1 -A3 "ELECT" $S
} > profiles/r2_sass_excerpt_headline.txt
wc -l profiles/r2_sass_excerpt_headline.txt
