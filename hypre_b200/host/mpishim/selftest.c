/* hypre_b200/host/mpishim/selftest.c -- TEST INFRASTRUCTURE (built by oracle/Makefile only): exercises the message-passing shim itself
 * (run by tests/test_ref_multirank.py under mpirun_shim and by tests/test_native_host.py through SHIM_DIR). Prints "SHIM_OK". */
#include "mpi.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "rank %d: check failed line %d: %s\n", rank, __LINE__, #c); MPI_Abort(MPI_COMM_WORLD, 3); } } while (0)

static void merge_max(void *in, void *inout, int *len, MPI_Datatype *dt)
{
   (void)dt;
   for (int i = 0; i < *len; i++) { if (((int *)in)[i] > ((int *)inout)[i]) { ((int *)inout)[i] = ((int *)in)[i]; } }
}

int main(int argc, char **argv)
{
   int rank, size;
   MPI_Init(&argc, &argv);
   MPI_Comm_rank(MPI_COMM_WORLD, &rank);
   MPI_Comm_size(MPI_COMM_WORLD, &size);

   /* ring of large non-blocking messages (bigger than a socket buffer), both directions at once */
   int n = 300000;
   double *sb = malloc(sizeof(double) * n), *rb = malloc(sizeof(double) * n);
   for (int i = 0; i < n; i++) { sb[i] = rank * 1e6 + i; }
   MPI_Request q[2];
   int right = (rank + 1) % size, left = (rank + size - 1) % size;
   MPI_Irecv(rb, n, MPI_DOUBLE, left, 7, MPI_COMM_WORLD, &q[0]);
   MPI_Isend(sb, n, MPI_DOUBLE, right, 7, MPI_COMM_WORLD, &q[1]);
   MPI_Status st[2];
   MPI_Waitall(2, q, st);
   CHECK(st[0].MPI_SOURCE == left && st[0].MPI_TAG == 7);
   int cnt; MPI_Get_count(&st[0], MPI_DOUBLE, &cnt); CHECK(cnt == n);
   for (int i = 0; i < n; i += 997) { CHECK(rb[i] == left * 1e6 + i); }

   /* tags, ANY_SOURCE, probe, ordering between a pair */
   if (size > 1)
   {
      if (rank != 0)
      {
         int v1 = 100 + rank, v2 = 200 + rank;
         MPI_Send(&v1, 1, MPI_INT, 0, 1, MPI_COMM_WORLD);
         MPI_Send(&v2, 1, MPI_INT, 0, 2, MPI_COMM_WORLD);
         MPI_Send(&v1, 1, MPI_INT, 0, 3, MPI_COMM_WORLD);
         MPI_Send(&v2, 1, MPI_INT, 0, 3, MPI_COMM_WORLD);
      }
      else
      {
         for (int k = 1; k < size; k++)
         {
            MPI_Status s; int v, c;
            MPI_Probe(MPI_ANY_SOURCE, 2, MPI_COMM_WORLD, &s);       /* tag 2 although tag 1 arrived first */
            MPI_Get_count(&s, MPI_INT, &c); CHECK(c == 1 && s.MPI_TAG == 2);
            MPI_Recv(&v, 1, MPI_INT, s.MPI_SOURCE, 2, MPI_COMM_WORLD, &s);
            CHECK(v == 200 + s.MPI_SOURCE);
         }
         for (int k = 1; k < size; k++)
         {
            int v, w; MPI_Status s;
            MPI_Recv(&v, 1, MPI_INT, k, 1, MPI_COMM_WORLD, MPI_STATUS_IGNORE); CHECK(v == 100 + k);
            MPI_Recv(&v, 1, MPI_INT, k, MPI_ANY_TAG, MPI_COMM_WORLD, &s);   /* same tag twice: order kept */
            MPI_Recv(&w, 1, MPI_INT, k, 3, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
            CHECK(s.MPI_TAG == 3 && v == 100 + k && w == 200 + k);
         }
         int flag; MPI_Iprobe(MPI_ANY_SOURCE, MPI_ANY_TAG, MPI_COMM_WORLD, &flag, MPI_STATUS_IGNORE); CHECK(!flag);
      }
   }

   /* collectives */
   double x = rank + 1.0, s = 0.0;
   MPI_Allreduce(&x, &s, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD); CHECK(s == size * (size + 1) / 2.0);
   int m = rank, mx = -1, mn = -1;
   MPI_Allreduce(&m, &mx, 1, MPI_INT, MPI_MAX, MPI_COMM_WORLD); CHECK(mx == size - 1);
   MPI_Reduce(&m, &mn, 1, MPI_INT, MPI_MIN, size - 1, MPI_COMM_WORLD); if (rank == size - 1) { CHECK(mn == 0); }
   int sc = 0; MPI_Scan(&m, &sc, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD); CHECK(sc == rank * (rank + 1) / 2);
   int *all = malloc(sizeof(int) * size); MPI_Allgather(&m, 1, MPI_INT, all, 1, MPI_INT, MPI_COMM_WORLD);
   for (int r = 0; r < size; r++) { CHECK(all[r] == r); }
   int *cnts = malloc(sizeof(int) * size), *dsp = malloc(sizeof(int) * size), tot = 0;
   for (int r = 0; r < size; r++) { cnts[r] = r + 1; dsp[r] = tot; tot += r + 1; }
   int *mine = malloc(sizeof(int) * (rank + 1)), *gv = malloc(sizeof(int) * tot);
   for (int i = 0; i <= rank; i++) { mine[i] = 10 * rank + i; }
   MPI_Allgatherv(mine, rank + 1, MPI_INT, gv, cnts, dsp, MPI_INT, MPI_COMM_WORLD);
   for (int r = 0; r < size; r++) { for (int i = 0; i <= r; i++) { CHECK(gv[dsp[r] + i] == 10 * r + i); } }
   int b = rank == size - 1 ? 4242 : 0; MPI_Bcast(&b, 1, MPI_INT, size - 1, MPI_COMM_WORLD); CHECK(b == 4242);
   int *a2a = malloc(sizeof(int) * size), *r2a = malloc(sizeof(int) * size);
   for (int r = 0; r < size; r++) { a2a[r] = 1000 * rank + r; }
   MPI_Alltoall(a2a, 1, MPI_INT, r2a, 1, MPI_INT, MPI_COMM_WORLD);
   for (int r = 0; r < size; r++) { CHECK(r2a[r] == 1000 * r + rank); }
   int *sv = malloc(sizeof(int) * size), got = -1;
   for (int r = 0; r < size; r++) { sv[r] = 7 * r; }
   MPI_Scatter(sv, 1, MPI_INT, &got, 1, MPI_INT, 0, MPI_COMM_WORLD); CHECK(got == 7 * rank);
   MPI_Op op; MPI_Op_create(merge_max, 1, &op);
   int um[2] = { rank, -rank }, ur[2];
   MPI_Allreduce(um, ur, 2, MPI_INT, op, MPI_COMM_WORLD); CHECK(ur[0] == size - 1 && ur[1] == 0);
   MPI_Op_free(&op);

   /* sub-communicators: split by parity, reversed order inside; create from a group */
   MPI_Comm half;
   MPI_Comm_split(MPI_COMM_WORLD, rank % 2, -rank, &half);
   int hr, hs; MPI_Comm_rank(half, &hr); MPI_Comm_size(half, &hs);
   CHECK(hs == (size + 1 - rank % 2) / 2);
   int hsum = 0; MPI_Allreduce(&rank, &hsum, 1, MPI_INT, MPI_SUM, half);
   int expect = 0; for (int r = rank % 2; r < size; r += 2) { expect += r; }
   CHECK(hsum == expect);
   int top = -1, mymax = rank; MPI_Allreduce(&mymax, &top, 1, MPI_INT, MPI_MAX, half);
   if (hr == 0) { CHECK(rank == top); }                 /* key = -rank: highest world rank first */
   MPI_Comm_free(&half);
   MPI_Group wg, sub; MPI_Comm only;
   MPI_Comm_group(MPI_COMM_WORLD, &wg);
   int first = 0; MPI_Group_incl(wg, 1, &first, &sub);
   MPI_Comm_create(MPI_COMM_WORLD, sub, &only);
   CHECK((rank == 0) == (only != MPI_COMM_NULL));
   MPI_Group_free(&wg); MPI_Group_free(&sub);

   /* derived datatype: every second double */
   MPI_Datatype strided;
   MPI_Type_vector(4, 1, 2, MPI_DOUBLE, &strided); MPI_Type_commit(&strided);
   double src[8] = { 0, 1, 2, 3, 4, 5, 6, 7 }, dst[4] = { -1, -1, -1, -1 };
   MPI_Request tq; MPI_Irecv(dst, 4, MPI_DOUBLE, rank, 9, MPI_COMM_WORLD, &tq);
   MPI_Send(src, 1, strided, rank, 9, MPI_COMM_WORLD);           /* to self */
   MPI_Wait(&tq, MPI_STATUS_IGNORE);
   CHECK(dst[0] == 0 && dst[1] == 2 && dst[2] == 4 && dst[3] == 6);
   MPI_Type_free(&strided);

   MPI_Barrier(MPI_COMM_WORLD);
   if (rank == 0) { printf("SHIM_OK %d ranks\n", size); }
   MPI_Finalize();
   return 0;
}
