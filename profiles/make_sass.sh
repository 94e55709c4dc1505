#!/bin/bash
# Per-kernel SASS listings of the frame's kernels (default template instances) from the built library -> profiles/sass/<kernel>.sass,
# plus the tensor / TMA mnemonic counts north_star asks for. Run from anywhere: paths are relative to this script.
#   profiles/make_sass.sh
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
LIB="$HERE/../synerfgine_b200/libsnb.so"
OUT="$HERE/sass"
mkdir -p "$OUT"
cuobjdump -sass "$LIB" 2>/dev/null > /tmp/snb_all.sass
# name -> substring of the mangled function name
declare -A K=(
  [k_nerf_raygen_pinhole]="13k_nerf_raygenILb0EE"
  [k_nerf_raygen_camera_extras]="13k_nerf_raygenILb1EE"
  [k_nerf_march_fused_F2_alt_12warps]="18k_nerf_march_fusedILi2ELb1ELi12ELb0EE"
  [k_nerf_march_fused_F2_stock_12warps]="18k_nerf_march_fusedILi2ELb0ELi12ELb0EE"
  [k_nerf_march_fused_F2_alt_12warps_lockstep]="18k_nerf_march_fusedILi2ELb1ELi12ELb1EE"
  [k_nerf_march_ws_F2_alt]="15k_nerf_march_wsILi2ELb1EE"
  [k_network_inference_tc_F2]="22k_network_inference_tcILi2EE"
  [k_write_normals]="15k_write_normals"
  [k_nerf_shadow_terms]="19k_nerf_shadow_terms"
  [k_nerf_shadow_apply]="19k_nerf_shadow_apply"
  [k_syn_raygen]="12k_syn_raygen"
  [k_raytrace]="10k_raytrace"
  [k_overlay]="9k_overlay"
  [k_accumulate_tonemap]="20k_accumulate_tonemap"
  [k_present]="9k_present"
  [k_alt_verify]="12k_alt_verify"
  [k_alt_begin_pass]="16k_alt_begin_pass"
  [k_grid_scales]="13k_grid_scales"
  [k_lock_begin]="12k_lock_begin"
  [k_lock_next]="11k_lock_next"
  [k_lbvh_topology]="15k_lbvh_topology"
  [k_bvh_fit]="9k_bvh_fit"
)
SUMMARY="$OUT/SUMMARY.txt"
echo "# SASS listings of $(basename "$LIB") ($(date -u +%F)); instruction counts and tensor / TMA / TMEM mnemonics per kernel" > "$SUMMARY"
printf "%-40s %8s %8s %8s %8s %8s %8s %8s\n" kernel instrs HMMA UTCHMMA LDTM UBLKCP UTCBAR LDG >> "$SUMMARY"
for name in "${!K[@]}"; do
  pat="${K[$name]}"
  awk -v pat="$pat" '/Function : /{keep=index($0,pat)>0} keep' /tmp/snb_all.sass | sed -E 's# +/\* 0x[0-9a-f]{16} \*/##' | grep -vE '^\s+/\* 0x[0-9a-f]{16} \*/$' > "$OUT/$name.sass"
  n=$(grep -cE '^\s+/\*[0-9a-f]{4,}\*/' "$OUT/$name.sass" || true)
  c() { grep -cE "^\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P[0-9T] )?$1" "$OUT/$name.sass" || true; }
  printf "%-40s %8s %8s %8s %8s %8s %8s %8s\n" "$name" "$n" "$(c HMMA)" "$(c UTCHMMA)" "$(c LDTM)" "$(c UBLKCP)" "$(c UTCBAR)" "$(c LDG)" >> "$SUMMARY"
done
sort -o "$SUMMARY" "$SUMMARY"
cat "$SUMMARY"
