/* fixture_io.h -- tiny named-array container for oracle fixtures (TEST INFRASTRUCTURE).
 * File = "D4BFIX1\0" then records: name[32] | dtype int32 | pad int32 | count int64 | data.
 * Reader: oracle/fixture.py. */
#ifndef FIXTURE_IO_H
#define FIXTURE_IO_H
#include <stdio.h>
#include <string.h>
#include <stdint.h>
#include "d4est_b200.h"
enum { FIX_F64 = 0, FIX_I32 = 1, FIX_I64 = 2, FIX_U8 = 3 };
static const int fix_size[4] = { 8, 4, 8, 1 };
static FILE *fix_open(const char *path)
{ FILE *f = fopen(path, "wb"); if (!f) { perror(path); return NULL; } fwrite("D4BFIX1", 1, 8, f); return f; }
static void fix_write(FILE *f, const char *name, int dtype, int64_t count, const void *data)
{
  char nm[32]; memset(nm, 0, sizeof nm); strncpy(nm, name, 31);
  int32_t hdr[2] = { dtype, 0 };
  fwrite(nm, 1, 32, f); fwrite(hdr, 4, 2, f); fwrite(&count, 8, 1, f);
  if (count > 0) fwrite(data, (size_t)fix_size[dtype], (size_t)count, f);
}
static const char *fix_prefix = "";   /* e.g. "L2." for level 2 of a multigrid fixture */
static void fix_write_p(FILE *f, const char *name, int dtype, int64_t count, const void *data)
{ char nm[64]; snprintf(nm, sizeof nm, "%s%s", fix_prefix, name); fix_write(f, nm, dtype, count, data); }
static void fix_write_desc(FILE *f, const d4b_mesh_desc_t *D)
{
  int32_t hdr[4] = { D->n_local, D->n_ghost, D->quad_type, D->n_peers };
  int64_t nh = D->hm_begin[D->n_local];
#define fix_write fix_write_p
  fix_write(f, "hdr", FIX_I32, 4, hdr);
  fix_write(f, "deg", FIX_I32, D->n_local + D->n_ghost, D->deg);
  fix_write(f, "node_off", FIX_I64, D->n_local + D->n_ghost + 1, D->node_off);
  fix_write(f, "degq", FIX_I32, D->n_local, D->degq);
  fix_write(f, "quad_off", FIX_I64, D->n_local + 1, D->quad_off);
  fix_write(f, "G", FIX_F64, 6 * D->quad_off[D->n_local], D->G);
  fix_write(f, "hm_begin", FIX_I32, D->n_local + 1, D->hm_begin);
  fix_write(f, "hm", FIX_U8, nh * (int64_t)sizeof(d4b_hm_t), D->hm);
  fix_write(f, "fac", FIX_F64, D->n_fac, D->fac);
  fix_write(f, "bc_g", FIX_F64, D->n_bc_g, D->bc_g);
  fix_write(f, "peer_rank", FIX_I32, D->n_peers, D->peer_rank);
  fix_write(f, "send_begin", FIX_I32, D->n_peers + 1, D->send_begin);
  fix_write(f, "send_elem", FIX_I32, D->send_begin[D->n_peers], D->send_elem);
  fix_write(f, "recv_begin", FIX_I32, D->n_peers + 1, D->recv_begin);
#undef fix_write
}
#endif
