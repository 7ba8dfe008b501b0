/* Test stand-in for the three BLACS / ScaLAPACK TOOLS symbols the reference's pdpotrf wrapper binds
 * (src/scalapack_wrappers/common.h:298-301).  Loaded RTLD_GLOBAL by the tests so that libdplasma_b200 finds
 * them in the process, the way it finds a real BLACS in an application linked with ScaLAPACK.
 * Only the contexts registered with fake_blacs_set answer; any other context reports "not in the grid". */
static int g_ctxt[8], g_P[8], g_Q[8], g_r[8], g_c[8], g_n = 0;
int g_fake_blacs_calls = 0;
void fake_blacs_set(int ctxt, int P, int Q, int r, int c) {
    int i;
    for (i = 0; i < g_n; i++) if (g_ctxt[i] == ctxt) break;
    if (i == g_n && g_n < 8) g_n++;
    if (i < 8) { g_ctxt[i] = ctxt; g_P[i] = P; g_Q[i] = Q; g_r[i] = r; g_c[i] = c; }
}
void Cblacs_gridinfo(int ctxt, int *P, int *Q, int *r, int *c) {
    g_fake_blacs_calls++;
    for (int i = 0; i < g_n; i++)
        if (g_ctxt[i] == ctxt) { *P = g_P[i]; *Q = g_Q[i]; *r = g_r[i]; *c = g_c[i]; return; }
    *P = *Q = *r = *c = -1;
}
void Cblacs_get(int ctxt, int what, int *val) { (void)ctxt; (void)what; *val = 0; }
int numroc_(int *n, int *nb, int *iproc, int *isrcproc, int *nprocs) {
    int mydist = (*nprocs + *iproc - *isrcproc) % *nprocs, nblocks = *n / *nb;
    int cnt = (nblocks / *nprocs) * *nb, extra = nblocks % *nprocs;
    if (mydist < extra) cnt += *nb; else if (mydist == extra) cnt += *n % *nb;
    return cnt;
}
