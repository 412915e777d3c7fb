for B in 148 296 444 592 1184; do
for LM in 0 100000; do
B200TFHE_LAT_MAX=$LM python bench.py --batch $B --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sw.json 2>> gpurun_out/sw.err
python -c "
import json; d=json.load(open('gpurun_out/sw.json')); print($B, $LM, round(d['value']), d['roofline']['ms_per_launch'], d['roofline']['kernel'])"
done; done
