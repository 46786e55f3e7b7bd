/* Appended (on gcc's stdin) behind the unmodified impl/himenoBMTxps.c by build_dropin.sh, so that the test can reach the
 * reference's file-scope arrays and run its own initmt() / jacobi() (impl/himenoBMTxps.c:206-299).  TEST INFRASTRUCTURE ONLY. */
float* href_array(int which) {
  switch (which) {
    case 0: return &p[0][0][0];
    case 1: return &a[0][0][0][0];
    case 2: return &b[0][0][0][0];
    case 3: return &c[0][0][0][0];
    case 4: return &bnd[0][0][0];
    case 5: return &wrk1[0][0][0];
    case 6: return &wrk2[0][0][0];
    default: return 0;
  }
}
void href_dims(int* d) { d[0] = MIMAX; d[1] = MJMAX; d[2] = MKMAX; d[3] = imax; d[4] = jmax; d[5] = kmax; }
/* the set-up of main() (:95-112) for one rank, then nn sweeps; returns gosa */
float href_run(int nn, float om, int do_init) {
  if (do_init) {
    int mx = MX0 - 1, my = MY0 - 1, mz = MZ0 - 1;
    ndx = NDX0; ndy = NDY0; ndz = NDZ0;
    npe = ndx * ndy * ndz; id = 0;
    initcomm(ndx, ndy, ndz);
    int it = initmax(mx, my, mz);
    initmt(mx, it);
  }
  omega = om;
  return jacobi(nn);
}
