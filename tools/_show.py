import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, 'ERR', e); continue
    print(f, 'ms/step %.3f value %.1f e2e %.1f launches %s roof %.3g/%.3g' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('gpu_launches'), d['roofline'].get('achieved') or 0, d['roofline'].get('frac') or 0))
    for k, v in sorted(d['profile_one_step'].items(), key=lambda kv: -kv[1]['ms'])[:7]:
        print('   %-22s %9.3f ms  %3d groups  %.3f' % (k, v['ms'], v['launch_groups'], v['share']))
