import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.3f M/s  ms/step %.4f  e2e %.3f M/s  step frac %.4f" % (d['value'] / 1e6, d['ms_per_step'], d['e2e']['value'] / 1e6, d['roofline_step']['frac']))
for k, v in d['kernels'].items():
    print("  %-22s %8.2f us  %s GB/s" % (k, v['avg_us'], v['GBps']))
