import json, sys
rows=[json.loads(l) for l in open(sys.argv[1])]
tot={}; roof={}
for r in rows:
    tot[r['kind']]=tot.get(r['kind'],0)+r['ms']; roof[r['kind']]=roof.get(r['kind'],0)+r['roof_ms']
    extra = f"S={r.get('tile_s')} NT={r.get('tile_n')} nA={r.get('a_stages')} nB={r.get('b_stages')} res={r.get('b_resident')}" if 'tile_s' in r else f"scratch={r.get('scratch_mb')}MB"
    print(f"{r['layer']:8s} {r['kind']:5s} {str(r['shape']):28s} {r['ms']:8.3f} ms {r['tflops']:7.1f} TF/s {r['gbs']:7.1f} GB/s roof {r['roof_ms']:.3f} frac {r['frac_of_roof_measured']:.2f} {extra}")
print({k:round(v,3) for k,v in tot.items()}, {k:round(v,3) for k,v in roof.items()})
