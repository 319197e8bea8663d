import sys, json
for line in sys.stdin:
    line = line.strip()
    if line.startswith('{'):
        d = json.loads(line)
        keep = {k: (round(v, 3) if isinstance(v, float) else v) for k, v in d.items()
                if k.endswith('ms') or k.endswith('tflops') or k in ('timeout', 'rc', 'stderr')}
        rel = {k: round(d[k]['rel'], 5) for k in ('fprop', 'dgrad', 'wgrad', 'err') if isinstance(d.get(k), dict)}
        print(d.get('case'), keep, rel)
    elif 'gpurun' in line or line.startswith('##'):
        print(line)
