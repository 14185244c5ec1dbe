#!/bin/bash
# SASS census of the shipped library: mnemonic counts per cubin (after a build): scratch/sass_census.sh > profiles/r2_sass_census.txt
set -e
lib=/root/repo/fit-slam_b200/libfitslam_cuda.so
tmp=$(mktemp -d); (cd $tmp && cuobjdump -xelf all $lib > /dev/null)
echo "# cuobjdump -sass of $(basename $lib) (built from $(git -C /root/repo rev-parse --short HEAD) + working tree), instruction counts per cubin"
printf "%-12s %8s %7s %8s %7s %7s %7s %7s %7s %7s %7s %7s %8s\n" cubin total UBLKCP UTMALDG SYNCS ATOMS ATOMG REDUX MUFU FP64 HMMA UTCMMA UCGABAR
for f in $tmp/*.cubin; do
  s=$(cuobjdump -sass $f)
  c() { echo "$s" | grep -cE "^\s+/\*[0-9a-f]{4}\*/\s+(@!?U?P[0-9T] )?($1)" || true; }
  printf "%-12s %8s %7s %8s %7s %7s %7s %7s %7s %7s %7s %7s %8s\n" $(basename $f .sm_100a.cubin | cut -c1-12) $(c "[A-Z]") $(c UBLKCP) $(c UTMALDG) $(c SYNCS) $(c ATOMS) $(c "ATOMG|ATOM\.") $(c REDUX) $(c MUFU) $(c "DADD|DMUL|DFMA") $(c HMMA) $(c "UTC.MMA") $(c UCGABAR)
done
rm -rf $tmp
