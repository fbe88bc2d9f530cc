# cold-launch policy: how many half-sweeps after a random initialisation use the pair-scored cost matrix
for k in 0 1 2 3 4 6; do
  echo "== ACMMP_COLD_SWEEPS=$k"
  ACMMP_COLD_SWEEPS=$k timeout 300 python tools/run_one_view.py --runs 3 2>&1 | tail -n 2 | head -n 1
  ACMMP_COLD_SWEEPS=$k timeout 300 python tools/scan_check.py --views 12 --width 1600 --height 1200 --nsrc 10 2>&1 | grep -E 'stage|^crc' | python -c "
import sys, json
tot = 0; sw = 0; out = []
crc = None
for l in sys.stdin:
    if l.startswith('crc'):
        crc = l.split()[1]; continue
    d = json.loads(l); tot += d['kernel_ms']; sw += d['sweeps_ms']; out.append('%s/%d:%.0f' % (d['stage'][:5], d['level'], d['kernel_ms']))
print('scan12 kernel_ms %.1f sweeps_ms %.1f  ' % (tot, sw), ' '.join(out), 'crc', crc)
"
done
