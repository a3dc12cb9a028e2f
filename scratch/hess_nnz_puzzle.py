"""Brute-force search for a block model of the docs-era Lagrangian Hessian nnz counts
(multilevel_transmon.jl:107: 11223; two_qubit_gates.jl:137: 22843)."""
import itertools
cases = [dict(N=50, x=50, m=2, sub=8, target=11223), dict(N=100, x=32, m=4, sub=32, target=22843)]
def tri(n): return n * (n + 1) // 2
opts = {
 'xu': [lambda c: c['x'] * c['m']],
 'xt': [lambda c: c['x'], lambda c: 0],
 'uu': [lambda c: tri(c['m']), lambda c: c['m'] ** 2, lambda c: c['m']],
 'ut': [lambda c: c['m'], lambda c: 0],
 'tt': [lambda c: 1, lambda c: 0],
 'xx': [lambda c: 0, lambda c: c['x'], lambda c: tri(c['x']), lambda c: c['x'] ** 2],
 'deriv': [lambda c: 2 * c['m'], lambda c: 0, lambda c: 4 * c['m']],
 'reg_knot': [lambda c: 3 * c['m'], lambda c: 6 * c['m'] + 1, lambda c: 6 * c['m'], lambda c: 3 * c['m'] + 1],
 'term': [lambda c: tri(c['x']), lambda c: c['x'] ** 2, lambda c: tri(c['sub']), lambda c: c['sub'] ** 2, lambda c: c['x']],
 'fix': [0, 1, 2],     # 0: nothing removed; 1: x_1 rows/cols removed; 2: x_1, u_1, u_N removed
 'reg_N': [0, 1],      # regularisers over N or N-1 knots
 'dyn_x2': [0, 1],     # dynamics Hessian also over x_{k+1}? (adds nothing) placeholder
}
keys = list(opts)
hits = []
for combo in itertools.product(*[range(len(opts[k])) for k in keys]):
    sel = dict(zip(keys, combo))
    ok = True
    for c in cases:
        g = lambda k: opts[k][sel[k]](c)
        K = c['N'] - 1
        dyn = g('xu') + g('xt') + g('uu') + g('ut') + g('tt') + g('xx')
        tot = K * dyn + K * g('deriv') + (c['N'] - sel['reg_N']) * g('reg_knot') + g('term')
        if sel['fix'] >= 1:   # remove entries touching x_1
            tot -= g('xu') + g('xt') + g('xx')
        if sel['fix'] >= 2:   # u_1 fixed: its uu, ut entries and reg entries; u_N fixed: reg entries only
            tot -= g('uu') + g('ut')
            if sel['fix'] >= 2 and g('xu') and sel['fix'] < 1: pass
        if tot != c['target']: ok = False; break
    if ok: hits.append(sel)
print(len(hits))
for h in hits[:20]: print(h)
