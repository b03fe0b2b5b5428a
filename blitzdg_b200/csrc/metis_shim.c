/* libbdg_metis.so -- host-only shim around the METIS 5.1.0 archive that ships with the CUDA toolkit
 * (libmetis_static.a, no header), for MeshManager::partitionMesh (reference src/MeshManager.cpp:491-544).
 *
 * Not part of the kernel library: setup-side host code, loaded by blitzdg_b200/mesh_manager.py.
 * The archive was built with 64-bit idx_t (probe: METIS_SetDefaultOptions fills 40 eight-byte slots with -1;
 * bdg_metis_idx_bytes() repeats the probe at run time), so the reference's int32 index_type arrays are widened
 * here.  Option slots follow metis.h 5.1.0 (the archive has Match_2Hop, i.e. METIS_OPTION_NO2HOP exists).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int64_t idx_t;

int METIS_SetDefaultOptions(idx_t *options);
int METIS_PartMeshNodal(idx_t *ne, idx_t *nn, idx_t *eptr, idx_t *eind, idx_t *vwgt, idx_t *vsize, idx_t *nparts,
                        void *tpwgts, idx_t *options, idx_t *objval, idx_t *epart, idx_t *npart);
int METIS_PartMeshDual(idx_t *ne, idx_t *nn, idx_t *eptr, idx_t *eind, idx_t *vwgt, idx_t *vsize, idx_t *ncommon,
                       idx_t *nparts, void *tpwgts, idx_t *options, idx_t *objval, idx_t *epart, idx_t *npart);

enum { /* metis.h 5.1.0 moptions_et */
  OPT_PTYPE = 0, OPT_OBJTYPE, OPT_CTYPE, OPT_IPTYPE, OPT_RTYPE, OPT_DBGLVL, OPT_NITER, OPT_NCUTS, OPT_SEED, OPT_NO2HOP,
  OPT_MINCONN, OPT_CONTIG, OPT_COMPRESS, OPT_CCORDER, OPT_PFACTOR, OPT_NSEPS, OPT_UFACTOR, OPT_NUMBERING, OPT_HELP,
  OPT_TPWGTS, OPT_NCOMMON, OPT_NOOUTPUT, OPT_BALANCE, OPT_GTYPE, OPT_UBVEC, NOPTIONS = 40
};

/* width of the archive's idx_t in bytes (8 expected), 0 if the probe is inconclusive */
int bdg_metis_idx_bytes(void) {
  int64_t buf[2 * NOPTIONS];
  memset(buf, 0x5a, sizeof(buf));
  METIS_SetDefaultOptions((idx_t *)buf);
  int n64 = 0, n32 = 0;
  const int32_t *b32 = (const int32_t *)buf;
  for (int i = 0; i < NOPTIONS; ++i) n64 += buf[i] == -1;
  for (int i = 0; i < NOPTIONS; ++i) n32 += b32[i] == -1;
  if (n64 == NOPTIONS && buf[NOPTIONS] != -1) return 8;
  if (n32 == NOPTIONS && b32[NOPTIONS] != -1) return 4;
  return 0;
}

/* MeshManager::partitionMesh: METIS_PartMeshNodal with the reference's option set (src/MeshManager.cpp:498-514):
 * k-way, edge-cut objective, SHEM coarsening, METISRB initial partitioning, greedy refinement, 1 cut, 10 iterations,
 * ufactor 30, 0-based, minconn, contiguous.  seed: the reference passes -1 (time-dependent); tests pass a fixed one.
 * dual != 0: METIS_PartMeshDual with ncommon = 2 (elements sharing an edge) instead -- not the reference's call, kept
 * for studies.  Returns the METIS status (1 = METIS_OK), -100 on a width mismatch, -101 out of memory. */
int bdg_metis_part_mesh(int32_t ne, int32_t nv, int32_t nfaces, const int32_t *etov, int32_t nparts, int32_t seed,
                        int32_t dual, int32_t verbose, int32_t *epart_out, int32_t *npart_out, int64_t *objval_out) {
  if (bdg_metis_idx_bytes() != 8) return -100;
  idx_t NE = ne, NV = nv, NP = nparts, objval = 0, ncommon = 2;
  idx_t *eptr = (idx_t *)malloc(sizeof(idx_t) * ((size_t)ne + 1));
  idx_t *eind = (idx_t *)malloc(sizeof(idx_t) * (size_t)ne * nfaces);
  idx_t *epart = (idx_t *)calloc((size_t)ne, sizeof(idx_t));
  idx_t *npart = (idx_t *)calloc((size_t)nv, sizeof(idx_t));
  if (!eptr || !eind || !epart || !npart) {
    free(eptr); free(eind); free(epart); free(npart);
    return -101;
  }
  for (int64_t i = 0; i <= ne; ++i) eptr[i] = (idx_t)nfaces * i; /* homogeneous element type (:518-522) */
  for (int64_t i = 0; i < (int64_t)ne * nfaces; ++i) eind[i] = etov[i];
  idx_t opt[NOPTIONS];
  METIS_SetDefaultOptions(opt);
  opt[OPT_PTYPE] = 1;     /* METIS_PTYPE_KWAY */
  opt[OPT_OBJTYPE] = 0;   /* METIS_OBJTYPE_CUT */
  opt[OPT_CTYPE] = 1;     /* METIS_CTYPE_SHEM */
  opt[OPT_IPTYPE] = 4;    /* METIS_IPTYPE_METISRB */
  opt[OPT_RTYPE] = 1;     /* METIS_RTYPE_GREEDY */
  opt[OPT_NCUTS] = 1;
  opt[OPT_NITER] = 10;
  opt[OPT_SEED] = seed;
  opt[OPT_UFACTOR] = 30;
  opt[OPT_NUMBERING] = 0;
  opt[OPT_DBGLVL] = verbose ? 1 : 0; /* the reference asks for METIS_DBG_INFO; silent unless verbose */
  opt[OPT_MINCONN] = 1;
  opt[OPT_NOOUTPUT] = 0;
  opt[OPT_CONTIG] = 1;
  int rc;
  if (nparts <= 1) {
    rc = 1; /* METIS rejects nparts = 1; the map is all zeros */
  } else if (dual) {
    rc = METIS_PartMeshDual(&NE, &NV, eptr, eind, NULL, NULL, &ncommon, &NP, NULL, opt, &objval, epart, npart);
  } else {
    rc = METIS_PartMeshNodal(&NE, &NV, eptr, eind, NULL, NULL, &NP, NULL, opt, &objval, epart, npart);
  }
  for (int64_t i = 0; i < ne; ++i) epart_out[i] = (int32_t)epart[i];
  for (int64_t i = 0; i < nv; ++i) npart_out[i] = (int32_t)npart[i];
  if (objval_out) *objval_out = objval;
  free(eptr); free(eind); free(epart); free(npart);
  return rc;
}
