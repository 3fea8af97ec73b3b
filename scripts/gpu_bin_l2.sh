# binning: L2 behaviour of the far-hit reductions for a few configurations (event time of 4 calls, then a light ncu pass)
P=gpurun_out/${1:-binl2}
mkdir -p gpurun_out
M=lts__t_sectors_srcunit_tex_op_red_lookup_miss.sum,lts__t_sectors_srcunit_tex_op_red_lookup_hit.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python scripts/bin_once.py C3 4 > ${P}_${name}.log 2>&1; echo "$name rc=$?: $(tail -1 ${P}_${name}.log | cut -c1-70)"
  env "$@" timeout 300 ncu --metrics $M --clock-control none -k regex:k_bin_band -c 1 --csv --log-file ${P}_${name}_ncu.csv python scripts/bin_once.py C3 1 > /dev/null 2>&1
  grep -E "k_bin_band" ${P}_${name}_ncu.csv | awk -F'","' '{print $(NF-2), $NF}' | tr -d '"' | tr '\n' ';'; echo
}
run base HC_BIN_ROW_EVICT_LAST=0
run rowlast HC_BIN_ROW_EVICT_LAST=1
run t512 HC_BIN_BAND_THREADS=512
run t512_rowlast HC_BIN_BAND_THREADS=512 HC_BIN_ROW_EVICT_LAST=1
run band256 HC_BIN_BAND_B=256
run band4096 HC_BIN_BAND_B=4096
