#!/bin/bash
# ncu --set full, one launch per kernel the configurations' numbers rest on (run under gpurun; reports
# land in gpurun_out/prof_r02/).  Each case first runs plainly, then under ncu.
set -u
OUT=gpurun_out/prof_r02
mkdir -p $OUT
run() {  # case, kernel regex, launches to skip, report name
  timeout 120 python tools/profile_driver.py $1 > $OUT/$4.plain.log 2>&1 || { echo "plain run of $1 failed"; return; }
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o $OUT/$4 \
      python tools/profile_driver.py $1 > $OUT/$4.ncu.log 2>&1
  # the report stays on the box (tens of MB each): its summary and the hottest SASS lines come back
  python tools/ncu_summary.py $OUT/$4.ncu-rep > $OUT/$4.json 2>> $OUT/$4.ncu.log
  python tools/ncu_hot.py $OUT/$4.ncu-rep 12 > $OUT/$4.hot.txt 2>> $OUT/$4.ncu.log
  rm -f $OUT/$4.ncu-rep
}
run headline       'scan_tma_kernel'        1 scan_tma_16_0_3_2_headline
run throughput     'scan_ldg_kernel'        1 scan_ldg_16_0_3_throughput
run c3_hbm         'scan_tma_kernel'        1 scan_tma_8_1_3_2_c3
run c3_throughput  'scan_ldg_kernel'        1 scan_ldg_8_1_3_c3
run c4             'scan_ldg_kernel'        1 scan_ldg_4_0_3_c4
run c5             'scan_'                  1 scan_c5_lsh256
run headline       'merge_sparse_kernel'    1 merge_sparse
run headline       'tau_init_hamming'       1 tau_init_hamming
run c3_hbm         'tau_init_euclid'        1 tau_init_euclid
run sig_lsh        'sig_lsh_kernel'         1 sig_lsh_16
run sig_lsh        'cache_fill_kernel'      0 cache_fill
run sig_minhash    'sig_minhash_kernel'     1 sig_minhash
ls -la $OUT
