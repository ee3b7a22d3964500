D=$PWD/medical-shape-disentanglement_b200/dsdf_b200
for lib in libdsdf_b200.base.so libdsdf_b200.so libdsdf_b200.base.so libdsdf_b200.so; do
DSDF_LIB=$D/$lib python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline --soak 300 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['sustained']; print('$lib', round(d['ms_per_step'],3), 'soak', round(s['ms_per_step'],3), s['clocks']['sm_mhz'], s['clocks']['power_w_max'])"
done
