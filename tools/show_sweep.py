import json, sys
lib = None
for l in open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/sweep_variants.jsonl'):
  try:
    d = json.loads(l)
  except Exception:
    continue
  if 'lib' in d:
    lib = d; continue
  print('%-26s %-24s value %10d  e2e %10d  kernel_ms %8.3f' % (lib['lib'][-26:], lib['spec'], d['value'], d['e2e']['value'], d['roofline']['kernel_ms']))
