#!/bin/bash
# r2zb: every BASELINE.json configuration (configs/*.json, the reference's input schema) through the reference-style CLI, final tree
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
cd dealii_multigrid_b200/lib
C=../../configs
timeout 900 ./multigrid_throughput $C/c1_hypercube_L6_p1.json $C/c2a_hypercube_L6_p4.json $C/c2b_hypercube_L7_p4.json \
  $C/c3_quadrant_L6_p1_global.json $C/c3_quadrant_L6_p1_local.json $C/c3_quadrant_L6_p4_global.json $C/c3_quadrant_L6_p4_local.json \
  $C/c4a_annulus_L6_p4_hpmg.json $C/c4b_annulus_L6_p4_pmg.json $C/c5_quadrant_L7_p4_gaussian.json > ../../$O/r2zb_configs_raw.txt 2>&1
echo "cli rc=$?"
tail -14 ../../$O/r2zb_configs_raw.txt | cut -c1-200
