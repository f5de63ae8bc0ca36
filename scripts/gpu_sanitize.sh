#!/bin/bash
# compute-sanitizer over the kernel protocols (mbarrier ring, named barrier, async-proxy fence, bulk stores, level
# counters through L2, last-ticket statistics): memcheck + racecheck + synccheck on the smoke shape (CTA teams), a
# one-CTA-per-world case, a tail-grid case and the 16-system set.  Summaries -> gpurun_out/sanitize_<tag>.txt
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() { # tag tool args...
    tag=$1; tool=$2; shift 2
    timeout 600 $CS --tool $tool --print-limit 20 python scripts/sanitize_case.py "$@" > gpurun_out/sanitize_${tag}_${tool}.txt 2>&1
    echo "$tag $tool rc=$? $(grep -c 'ERROR SUMMARY\|RACECHECK SUMMARY' gpurun_out/sanitize_${tag}_${tool}.txt) $(grep 'SUMMARY' gpurun_out/sanitize_${tag}_${tool}.txt | tr '\n' ' ') $(grep parity gpurun_out/sanitize_${tag}_${tool}.txt | cut -c1-120)"
}
for tool in memcheck racecheck synccheck; do
    run smoke_teams $tool chain3 16 1024 4 2
    run one_cta_per_world $tool chain3 24 1024 4 2 FIMO_B200_TEAM=1
    run tail_grid $tool dag8 1250 100 4 2 FIMO_B200_TEAM=1
    run dag16_set2 $tool dag16 40 512 4 2 FIMO_B200_TEAM=1
done
