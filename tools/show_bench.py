import json,sys
for path in sys.argv[1:]:
    d=json.load(open(path))
    print('===',path, {k:(round(d[k],1) if isinstance(d[k],float) else d[k]) for k in ['value','ms_per_step','queries_per_sec','gpu_launches']}, 'whole_frac', d['roofline']['whole_step']['frac'] if d.get('roofline') else None)
    for k in d['kernels']: print('   %-20s launches=%5d ms=%9.3f alg_gb=%9.3f alg_gbs=%s'%(k['name'],k['launches'],k['ms'],k['alg_gb'],k['alg_gbs']))
