"""Dev tool: the numbers of a bench.py JSON line that matter, one per line."""
import json, sys
for line in open(sys.argv[1]):
    if not line.startswith('{'):
        continue
    d = json.loads(line)
    print('value %.0f Mrays/s  ms_per_step %.4f  n_gpus %d  launch %s' % (d['value'], d['ms_per_step'], d['n_gpus'], d['config'].get('launch')))
    print('collective:', d['config'].get('collective'))
    r = d.get('roofline', {})
    if 'backward' in r:
        print('fwd %.4f ms frac %.3f | bwd %.4f ms frac %.3f' % (r['kernel_ms'], r['frac'], r['backward']['kernel_ms'], r['backward']['frac']))
    e = d.get('extras', {})
    for k in ('nccl_allreduce_step', 'peer_scatter_step'):
        if e.get(k): print(k, {a: b for a, b in e[k].items() if a != 'what'})
    if 'full_step_ms' in e:
        print('isect %.4f  full_step %.4f (%s)  shardcheck %s  gradsum %.6e' % (e['intersect_only_ms'], e['full_step_ms'], e.get('full_step_launch'), e.get('grad_rel_l2_vs_single_rank'), e['grad_abs_sum']))
        rs = e['render_step']; print('render step %.4f (%s; %s) kernel %.4f' % (rs['ms_per_step'], rs['launch'], rs.get('collective'), rs['roofline']['kernel_ms']), rs.get('e2e'))
    print('e2e', d.get('e2e')); print('cpu', d.get('cpu_baseline')); print('clocks', d.get('clocks'))
