import json, sys
d = json.load(open(sys.argv[1]))
print('value %.1f M/s  e2e %.1f M/s  kernel_ms %.4f  frac %.3f  tail_ms %.4f  ms/step %.3f' % (d['value'] / 1e6, d['e2e']['value'] / 1e6, d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['tail_kernel_ms'], d['ms_per_step']))
