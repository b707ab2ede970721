set -x
cd tools
for lim in -1 32 64 128; do ./gather_gran $lim 0; done
for m in 1 2 3 4 5; do ./gather_gran -1 $m; ./gather_gran 32 $m; done
./gather_gran -1 0 56; ./gather_gran 32 0 56
M="dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_lookup_miss.sum,lts__t_requests_srcunit_tex_op_read.sum"
for cfg in "-1 0" "32 0" "64 0" "-1 1" "32 1" "-1 2" "32 2" "-1 3" "-1 5" "32 5"; do
  ./gather_gran $cfg > /dev/null && ncu --metrics $M --clock-control none -s 1 -c 1 --csv ./gather_gran $cfg 2>&1 | grep -E "dram__bytes_read|tex_op_read|lookup_miss|limit_req" | sed "s/^/[$cfg] /"
done
cd ..
