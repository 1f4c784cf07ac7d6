"""Generates lassim_b200/csrc/rk_steppers.cuh: one fully unrolled step attempt
per explicit Runge-Kutta pair (Dormand-Prince 5(4) and Dormand-Prince 8(5,3)).

Why generated: the hot loop wants every coefficient at a fixed place (a slot of
one __constant__ array, read as an operand through a uniform register), zero
coefficients skipped, and the terms of the next stage state that do not depend
on the logistic being evaluated folded in advance (so that the dependent chain
from the logistic to the next stage state is a single FMA).  Writing that out by
hand for the 12-stage tableau invites typos; the 8(5,3) coefficients are taken
from SciPy's table (scipy.integrate._ivp.dop853_coefficients, which reproduces
Hairer's dop853.f) when the header is generated.

Usage: python tools/gen_rk_steppers.py   (rewrites the header; commit it)
"""
import os
from fractions import Fraction

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "lassim_b200", "csrc", "rk_steppers.cuh")


def lit(v):
    v = float(v)
    s = repr(v)
    if "e" not in s and "." not in s and "inf" not in s:
        s += ".0"
    return s


# A coefficient written as a literal costs two UMOV each time it is used (10-12 % of the executed
# instructions of the explicit kernels, `ncu` source page): with RK_CONST_POOL (default on) they
# live in a __constant__ array and arrive by one LDCU.64 each.  Measured: 16 Th2 populations 59.6 ->
# 58.2 ms, config 4 (1024 candidates) 2499 -> 2437 ms, the lone limit cycles unchanged.  The same
# change made the one-thread-per-candidate peripheral march 23 % SLOWER (8.4 -> 10.3 ms: nothing
# there hides the load's latency), so dop853_dense_tables.cuh keeps its literals.
USE_POOL = os.environ.get("RK_CONST_POOL", "1") != "0"
POOL = []


def con(v):
    if not USE_POOL:
        return lit(v)
    v = float(v)
    if v not in POOL:
        POOL.append(v)
    return "c_rk[%d]" % POOL.index(v)


def dp45():
    F = Fraction
    A = [[], [F(1, 5)], [F(3, 40), F(9, 40)], [F(44, 45), F(-56, 15), F(32, 9)],
         [F(19372, 6561), F(-25360, 2187), F(64448, 6561), F(-212, 729)],
         [F(9017, 3168), F(-355, 33), F(46732, 5247), F(49, 176), F(-5103, 18656)]]
    B = [F(35, 384), 0, F(500, 1113), F(125, 192), F(-2187, 6784), F(11, 84)]
    E = [F(71, 57600), 0, F(-71, 16695), F(71, 1920), F(-17253, 339200), F(22, 525), F(-1, 40)]
    A = [[float(x) for x in row] for row in A]
    C = [0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
    return dict(name="dp45", S=6, order=5, A=A, B=[float(b) for b in B], C=C,
                E=[[float(e) for e in E]])


def dop853():
    from scipy.integrate._ivp import dop853_coefficients as dc
    S = 12
    A = [[float(dc.A[s, j]) for j in range(s)] for s in range(S)]
    return dict(name="dop853", S=S, order=8, A=A, B=[float(b) for b in dc.B],
                C=[float(c) for c in dc.C[:S]] + [1.0],
                E=[[float(e) for e in dc.E5], [float(e) for e in dc.E3]])


def weighted_sum(coefs, names, indent):
    """Nested-fma expression for sum_j coefs[j] * names[j][r] (zeros skipped)."""
    terms = [(c, n) for c, n in zip(coefs, names) if c != 0.0]
    if not terms:
        return "0.0"
    expr = "%s * %s[r]" % (con(terms[0][0]), terms[0][1])
    for c, n in terms[1:]:
        expr = "fma(%s, %s[r], %s)" % (con(c), n, expr)
    return expr


def emit(tb):
    S, A, B, C, E = tb["S"], tb["A"], tb["B"], tb["C"], tb["E"]
    name = tb["name"]
    K = ["k%d" % (i + 1) for i in range(S + 1)]          # k1 .. k_{S+1}
    L = []
    w = L.append
    w("// %s: one step attempt of size hs from (y, k1)." % name)
    w("//   sig(ys, sr, se)  evaluates the logistic terms sigmoid(W ys) = sr + sr*se for this lane's")
    w("//                genes (reciprocal seed and Newton residual, see logistic_fast_parts)")
    w("//   gmax(x)      max of a non-negative float over the lanes of the group")
    w("//   s0, s1       high words of the logistic values at y (in) and at yn (out): the swing of")
    w("//                the NEXT attempt starts from the state this one ends in, so that a switch")
    w("//                before the first inner stage (c = %s) is seen as well" % ("%.3f" % C[1]))
    w("//   swing        jump guard (>= 2 switches it off).  The embedded estimates are blind to a")
    w("//                logistic term that switches early in the step (the stages there carry no")
    w("//                weight in the solution or its embedded companions): a term that swings by")
    w("//                more than this between two stages of the attempt is charged as the jump it")
    w("//                may be, error <= hs * vmax * swing, so that the step shrinks until the switch")
    w("//                is resolved - or, for a true discontinuity, located to the tolerance")
    w("// Outputs: yn (the order-%d solution), kn (first stage derivative of the next step);"
      % tb["order"])
    w("// returns the group's error measure (accept if <= 1), +inf when anything was non-finite.")
    w("//")
    w("// IF = true: integrating factor for exponentially GROWING genes.  With lambda = lp - gr")
    w("// (lp >= 0, gr >= 0 the growth rate of a knocked-down gene whose lambda went negative),")
    w("// the pair is applied to u(tau) = exp(-gr tau) y(t_n + tau), for which")
    w("//     u' = exp(-gr tau) G(exp(gr tau) u),   G(y) = -lp*y + vmax*sigmoid(W y),")
    w("// so the homogeneous growth is exact and the step size is set by the rest of the")
    w("// dynamics.  `lam` then holds lp, `k1`/`kn` are G(y)/G(yn), stage derivatives live in")
    w("// u-space (K_s = -lp*U_s + vmax*exp(-gr c_s h)*sg), errors are mapped back to y.  Genes")
    w("// with gr = 0 see exactly the plain formulas.")
    w("template <bool IF, int R, class Sig, class GMax>")
    w("__device__ __forceinline__ float %s_attempt(Sig &&sig, GMax &&gmax," % name)
    w("        const double (&lam)[R], const double (&vm)[R], const double (&gr)[R],")
    w("        const double *tab, const double (&y)[R],")
    w("        const double (&k1)[R], const double hs, const double rtol, const double atol,")
    w("        const float swing, const int (&s0)[R], int (&s1)[R], double (&yn)[R], double (&kn)[R]) {")
    decl = ", ".join("%s[R]" % k for k in K[1:S])
    w("    double %s;" % decl)
    w("    double ys[R], sg[R], sr[R], se[R], q[R], ly[R], cv[R], ef[R], ei[R], yz[R];")
    w("    // smallest / largest logistic value over the stages, as the high words of the doubles")
    w("    // (monotonic for positive values; integer min/max keep this off the FP64 pipe)")
    w("    int slo[R], shi[R];")
    w("#pragma unroll")
    w("    for (int r = 0; r < R; ++r) { ys[r] = fma(hs * %s, k1[r], y[r]); ef[r] = 1.0; ei[r] = 1.0;" % con(A[1][0]))
    w("                                  slo[r] = s0[r]; shi[r] = s0[r]; }")
    for s in range(2, S + 1):
        row = A[s] if s < S else B
        last = row[s - 1]
        c = C[s - 1]
        target = "ys" if s < S else "yn"
        w("    {   // sigma at stage %d's state (c = %s) -> %s, %s" % (s, lit(c), K[s - 1], "state of stage %d" % (s + 1) if s < S else "solution"))
        w("        const double hl = hs * %s;" % con(last))
        w("        if constexpr (IF) {")
        w("#pragma unroll")
        w("            for (int r = 0; r < R; ++r) {")
        w("                ef[r] = exp2_tab(gr[r] * (hs * %s), tab);" % con(c * 1.4426950408889634))
        w("                ei[r] = (gr[r] == 0.0) ? 1.0 : fast_rcp(ef[r]);")
        w("                yz[r] = ef[r] * ys[r];")
        w("            }")
        w("        }")
        w("#pragma unroll")
        w("        for (int r = 0; r < R; ++r) {")
        w("            ly[r] = -lam[r] * ys[r];")
        w("            cv[r] = IF ? (hl * vm[r]) * ei[r] : hl * vm[r];")
        w("            q[r] = fma(hl, ly[r], fma(hs, %s, y[r]));" % weighted_sum(row[:s - 1], K[:s - 1], 16))
        w("        }")
        w("        if constexpr (IF) sig(yz, sr, se); else sig(ys, sr, se);")
        w("#pragma unroll")
        w("        for (int r = 0; r < R; ++r) {")
        w("            // cv*sigma + q with the Newton step of the reciprocal folded in")
        w("            %s[r] = fma(cv[r] * sr[r], se[r], fma(cv[r], sr[r], q[r]));" % target)
        w("            sg[r] = fma(sr[r], se[r], sr[r]);")
        w("            slo[r] = min(slo[r], __double2hiint(sg[r])); shi[r] = max(shi[r], __double2hiint(sg[r]));")
        w("            %s[r] = fma(IF ? vm[r] * ei[r] : vm[r], sg[r], ly[r]);" % K[s - 1])
        w("        }")
        w("    }")
    assert abs(C[S - 1] - 1.0) < 1e-15, "the last stage sits at c = 1: ef is exp(gr*hs) there"
    # last evaluation f(yn) and the error estimates
    nE = len(E)
    early = all(E[m][S] == 0.0 for m in range(nE))     # the estimate does not need f(yn)
    w("    // back to y (ef = exp(gr*hs) from the last stage, at c = 1), then f(yn) (first stage of")
    w("    // the next step) and the error estimate%s" % ("s" if nE > 1 else ""))
    w("    double isc[R];")
    for m in range(nE):
        w("    double eq%d[R], ce%d[R];" % (m, m))
    w("#pragma unroll")
    w("    for (int r = 0; r < R; ++r) {")
    w("        if constexpr (IF) yn[r] *= ef[r];")
    w("        ly[r] = -lam[r] * yn[r];")
    for m in range(nE):
        e_last = E[m][S]
        base = "hs * (%s)" % weighted_sum(E[m][:S], K[:S], 8)
        w("        { const double base = %s;" % base)
        w("          const double eb = IF ? ef[r] * base : base;      // u-space sum mapped to y")
        if e_last != 0.0:
            w("          ce%d[r] = (hs * %s) * vm[r];" % (m, con(e_last)))
            w("          eq%d[r] = fma(hs * %s, ly[r], eb); }" % (m, con(e_last)))
        else:
            w("          ce%d[r] = 0.0;" % m)
            w("          eq%d[r] = eb; }" % m)
    w("        // error weight rtol*max(|y|,|yn|) + atol; its reciprocal to 1e-12 is plenty for a")
    w("        // step-size decision and keeps an IEEE division off the path")
    w("        isc[r] = fast_rcp(fma(rtol, fmax(fabs(y[r]), fabs(yn[r])), atol));")
    w("    }")

    def emit_measure():
        for m in range(nE):
            w("    double qd%d = 0.0;" % m)
        w("#pragma unroll")
        w("    for (int r = 0; r < R; ++r) {")
        w("        // jump guard: |error| <= hs * vmax * swing of a term that switched inside the step")
        w("        const double sw = __hiloint2double(shi[r], 0) - __hiloint2double(min(slo[r], shi[r]), 0);")
        w("        if (sw > (double)swing) qd0 = fmax(qd0, (fabs(hs) * sw) * (IF ? vm[r] * ef[r] : vm[r]) * isc[r]);")
        for m in range(nE):
            if E[m][S] != 0.0:
                w("        const double e%d = fma(ce%d[r], sg[r], eq%d[r]);" % (m, m, m))
            else:
                w("        const double e%d = eq%d[r];" % (m, m))
            w("        qd%d = fmax(qd%d, fabs(e%d) * isc[r]);      // fmax drops NaN: checked next" % (m, m, m))
            w("        if (e%d != e%d) qd%d = CUDART_INF;" % (m, m, m))
        w("    }")
        for m in range(nE):
            w("    const float n%d = gmax((qd%d <= 3.0e38) ? (float)qd%d : CUDART_INF_F);" % (m, m, m))

    if early:
        w("    // the estimate needs k1..k%d only: reduce it over the group BEFORE f(yn), so that the" % S)
        w("    // shuffles of the reduction overlap the dependent chain of that last evaluation")
        emit_measure()
        w("    sig(yn, sr, se);")
        w("#pragma unroll")
        w("    for (int r = 0; r < R; ++r) {")
        w("        kn[r] = fma(vm[r] * sr[r], se[r], fma(vm[r], sr[r], ly[r]));")
        w("        sg[r] = fma(sr[r], se[r], sr[r]);")
        w("        s1[r] = __double2hiint(sg[r]);")
        w("    }")
    else:
        w("    sig(yn, sr, se);")
        w("#pragma unroll")
        w("    for (int r = 0; r < R; ++r) {")
        w("        kn[r] = fma(vm[r] * sr[r], se[r], fma(vm[r], sr[r], ly[r]));")
        w("        sg[r] = fma(sr[r], se[r], sr[r]);")
        w("        s1[r] = __double2hiint(sg[r]);")
        w("    }")
        emit_measure()
    if nE == 1:
        w("    return n0;")
    else:
        w("    // Hairer's 8(5,3) measure err5^2 / sqrt(err5^2 + 0.01 err3^2), in the max-norm")
        w("    if (!(n0 < 1.0e18f) || !(n1 < 1.0e18f)) return CUDART_INF_F;")
        w("    const float den = fmaf(n0, n0, 0.01f * n1 * n1);")
        w("    return den > 0.0f ? n0 * n0 * rsqrtf(den) : 0.0f;")
    w("}")
    w("")
    return "\n".join(L)


def main(out=OUT):
    parts = ["// GENERATED by tools/gen_rk_steppers.py - do not edit by hand.",
             "// Fully unrolled step attempts of the explicit Runge-Kutta pairs used by",
             "// grn_cost_kernel.  Every stage state after the second is formed as",
             "//   ys = c*sigma + q,  c = hs*a_last*vm,  q = y + hs*sum_{older k} + hs*a_last*(-lambda*ys_prev)",
             "// evaluated as fma(c*sr, se, fma(c, sr, q)) with sigma = sr + sr*se (seed and Newton residual",
             "// of the reciprocal): two dependent operations from the seed to the next stage state.",
             "#pragma once", "#include <math_constants.h>", "", "namespace lassim {", "",
             "// 1/x for x > 0 from the SFU seed (MUFU.RCP64H) and one Newton step: relative error <= 1e-12",
             "__device__ __forceinline__ double fast_rcp(double x) {",
             "    double r;",
             "    asm(\"rcp.approx.ftz.f64 %0, %1;\" : \"=d\"(r) : \"d\"(x));",
             "    return fma(r, fma(-x, r, 1.0), r);",
             "}", "",
             "// 2^a for a >= 0 (clamped at 1000) with the kernel's 2^(j/1024) table in shared memory:",
             "// a = n + j/1024 + f, |f| <= 2^-11, degree-2 Taylor of (2^f - 1)/f (5e-16); exactly 1 for a = 0.",
             "__device__ __forceinline__ double exp2_tab(double a, const double *tab) {",
             "    a = fmin(a, 1000.0);",
             "    const double s = a + kExpMagic;",
             "    const int ki = __double2loint(s);",
             "    const double f = a - (s - kExpMagic);",
             "    const double tj = tab[ki & (kExpTable - 1)];",
             "    const double p = fma(f, fma(0.05550410866482158, f, 0.24022650695910072), 0.6931471805599453);",
             "    const double v = fma(tj, f * p, tj);",
             "    return __hiloint2double(__double2hiint(v) + ((ki >> kExpTableLog2) << 20), __double2loint(v));",
             "}", ""]
    bodies = [emit(dp45()), emit(dop853())]
    if POOL:
        rows = [", ".join(lit(v) for v in POOL[i:i + 4]) for i in range(0, len(POOL), 4)]
        parts.append("// tableau coefficients of both pairs (operands c[bank][offset] of the FP64 instructions)")
        parts.append("__constant__ double c_rk[%d] = {\n    %s};\n" % (len(POOL), ",\n    ".join(rows)))
    parts.extend(bodies)
    parts.append("}  // namespace lassim\n")
    with open(out, "w") as fh:
        fh.write("\n".join(parts))
    print("wrote", out)


if __name__ == "__main__":
    main()
