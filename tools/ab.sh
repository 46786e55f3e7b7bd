#!/bin/bash
# A/B on the GPU box: parity tests, then encode/decode times for CT x field (device resident, no CPU baseline)
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for f in smooth noisy constant; do for ct in ${CTS:-5 7}; do
  python bench.py --ct $ct --field $f --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$f ct=$ct value=%.0f GB/s enc=%.3f ms dec=%.3f ms ratio=%.2f frac_enc=%.3f' % (d['value'], c['compress_ms'], c['decompress_ms'], c['compression_ratio'], d['roofline']['compress']['frac']))"
done; done
