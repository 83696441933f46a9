for cost in "256,67,134,90,70,50,40" "256,67,134,120,100,70,50" "256,67,134,140,115,80,60"; do
 for sb in 32 64; do
  DVF_GRAM_COST=$cost DVF_GRAM_SB=$sb DVF_GRAM_DEBUG=1 timeout 200 python tools/gram_tune.py 1000000 500 30 2>&1 | tail -2
 done
done
for g in 10 11 12 13; do DVF_GRAM_COST="256,67,134,120,100,70,50" DVF_GRAM_SB=32 DVF_GRAM_G=$g DVF_GRAM_DEBUG=1 timeout 200 python tools/gram_tune.py 1000000 500 30 2>&1 | tail -2; done
for sb in 16 32 64; do for lag in 0 1; do
  echo "== ncu SB=$sb LAG=$lag"
  DVF_GRAM_SB=$sb DVF_GRAM_LAG=$lag DVF_GRAM_NOCHECK=1 timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum -k regex:^k_gram$ --launch-skip 3 -c 1 python tools/gram_tune.py 1000000 500 30 2>&1 | grep -E "dram__|gpu__time|lts__"
done; done
