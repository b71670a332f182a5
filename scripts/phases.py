import sys, json
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): continue
    d=json.loads(line)
    r=d['roofline']
    print("ms/step %.1f dev %.1f onesweep/launch %.2f | %s" % (d['ms_per_step'], r['whole_path']['device_ms_per_step'], r['ms_per_launch'], {k:round(v,1) for k,v in r['phases_ms'].items()}))
