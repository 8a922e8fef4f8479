// Self-test of the MPI stand-in: run as N processes (TCLB_MPI_RANK/TCLB_MPI_SIZE/TCLB_MPI_ID set by the caller).
// Exercises what the reference uses: Allreduce MIN/MAX/SUM/LOR, Bcast, Reduce, Gather(v)/Scatterv, and the halo pattern
// (Irecv for every direction, Isend, Waitany loop, Waitall) with messages larger than one ring and several messages
// per pair of ranks with different tags posted in a different order than they are sent.
#include <mpi.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CHECK(c) do { if (!(c)) { fprintf(stderr, "[%d] FAILED: %s (line %d)\n", rank, #c, __LINE__); MPI_Abort(MPI_COMM_WORLD, 1); } } while (0)
int main(int argc, char **argv) {
    MPI_Init(&argc, &argv);
    int rank, size;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    int v = rank + 1, r = 0;
    MPI_Allreduce(&v, &r, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD); CHECK(r == size * (size + 1) / 2);
    MPI_Allreduce(&v, &r, 1, MPI_INT, MPI_MAX, MPI_COMM_WORLD); CHECK(r == size);
    MPI_Allreduce(&v, &r, 1, MPI_INT, MPI_MIN, MPI_COMM_WORLD); CHECK(r == 1);
    int flag = rank == size - 1, any = 0;
    MPI_Allreduce(&flag, &any, 1, MPI_INT, MPI_LOR, MPI_COMM_WORLD); CHECK(any == 1);
    double d[3] = {1.0 * rank, 2.0, -1.0 * rank}, ds[3] = {0, 0, 0};
    MPI_Reduce(d, ds, 3, MPI_DOUBLE, MPI_SUM, 0, MPI_COMM_WORLD);
    if (rank == 0) CHECK(ds[0] == size * (size - 1) / 2.0 && ds[1] == 2.0 * size);
    char name[16] = "";
    if (rank == size - 1) snprintf(name, sizeof(name), "hello");
    MPI_Bcast(name, 16, MPI_CHAR, size - 1, MPI_COMM_WORLD); CHECK(name[0] == 'h' && name[4] == 'o');
    // gatherv / scatterv of rank-dependent pieces
    std::vector<int> counts(size), displs(size);
    int tot = 0;
    for (int i = 0; i < size; i++) { counts[i] = 1000 * (i + 1); displs[i] = tot; tot += counts[i]; }
    std::vector<double> mine(counts[rank], (double)rank), all(rank == 0 ? tot : 1);
    MPI_Gatherv(mine.data(), counts[rank], MPI_DOUBLE, all.data(), counts.data(), displs.data(), MPI_DOUBLE, 0, MPI_COMM_WORLD);
    if (rank == 0) for (int i = 0; i < size; i++) CHECK(all[displs[i]] == i && all[displs[i] + counts[i] - 1] == i);
    std::vector<double> back(counts[rank], -1.0);
    MPI_Scatterv(all.data(), counts.data(), displs.data(), MPI_DOUBLE, back.data(), counts[rank], MPI_DOUBLE, 0, MPI_COMM_WORLD);
    CHECK(back[0] == rank && back[counts[rank] - 1] == rank);
    // halo pattern: 4 messages to each of two neighbours, sizes up to 20 MB (> one 8 MB ring), 6 rounds
    if (size > 1) {
        const int up = (rank + 1) % size, dn = (rank + size - 1) % size;
        const size_t sz[4] = {12, 20u << 20, 4096, 3u << 20};
        for (int round = 0; round < 6; round++) {
            std::vector<std::vector<unsigned char>> out(8), in(8);
            MPI_Request rr[8], sr[8];
            for (int i = 0; i < 8; i++) {
                out[i].resize(sz[i % 4]); in[i].resize(sz[i % 4]);
                for (size_t k = 0; k < out[i].size(); k += 997) out[i][k] = (unsigned char)(rank * 31 + i * 7 + round + k);
            }
            for (int i = 7; i >= 0; i--) MPI_Irecv(in[i].data(), (int)in[i].size(), MPI_BYTE, i < 4 ? dn : up, i + 100 * round, MPI_COMM_WORLD, &rr[i]);
            for (int i = 0; i < 8; i++) MPI_Isend(out[i].data(), (int)out[i].size(), MPI_BYTE, i < 4 ? up : dn, i + 100 * round, MPI_COMM_WORLD, &sr[i]);
            for (int j = 0; j < 8; j++) {
                int i;
                MPI_Waitany(8, rr, &i, MPI_STATUSES_IGNORE);
                const int from = i < 4 ? dn : up;
                for (size_t k = 0; k < in[i].size(); k += 997) CHECK(in[i][k] == (unsigned char)(from * 31 + i * 7 + round + k));
            }
            MPI_Waitall(8, sr, MPI_STATUSES_IGNORE);
        }
    }
    MPI_Barrier(MPI_COMM_WORLD);
    MPI_Finalize();
    if (rank == 0) printf("mpi_stub selftest ok on %d ranks\n", size);
    return 0;
}
