import sys, itertools
sys.path.insert(0, "tests")
import numpy as np
import oracle_lib as O
import problems as P
from gpu_lib import binding as B
for m in (160, 128):
    n = m * m
    u = P.grid2d_state(m)
    J = P.grid2d_jacobian(u, m)
    M = P.identity_mass(n)
    b = np.full(n, 1.24)
    csc = O.assemble_jb(O.port, n, J, M, 50.0)
    for blocked, dmma, fused, mind in [(0,1,1,96),(1,0,1,96),(1,1,1,96),(1,1,0,96),(1,0,0,96),(1,1,1,8),(1,1,1,200)]:
        c = B.Context(n)
        c.set_option("chain", 0); c.set_option("bf_dmma", dmma); c.set_option("blocked_lu", blocked); c.set_option("bf_fused_panel", fused); c.set_option("blocked_lu_min_dim", mind)
        c.set_mass(M); c.set_jacobian(J)
        c.refactor(50.0)
        c.set_rhs(b)
        x = c.solve()
        eta, _ = O.backward_error(n, csc, x, b)
        print(m, "blocked", blocked, "dmma", dmma, "fused", fused, "min_dim", mind, "eta", eta, "max_front", c.stat("max_front"), "nbig", c.stat("n_big_fronts"), flush=True)
        c.close()
