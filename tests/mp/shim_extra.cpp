// A small MPI program in the style of the reference's tutorials that uses the shim's "close
// relatives" (Sendrecv, Test, Waitany, Iprobe, Allgather, Comm_split/dup, Dims_create, Cart_get,
// hvector, resized, Get_address, logical reductions).  Compiled by tests/test_reference_conformance.py
// against include/b200comm/shim/mpi.h and run with 4 ranks; prints SHIM_EXTRA PASSED on rank 0.
#include <mpi.h>

#include <cstdio>
#include <vector>

#define CHECK(cond)                                                           \
    do {                                                                      \
        if (!(cond)) { std::printf("rank %d: FAILED %s (line %d)\n", rank, #cond, __LINE__); ++fails; } \
    } while (0)

int main(int argc, char** argv) {
    int rank = -1, size = 0, fails = 0, flag = 0;
    MPI_Initialized(&flag);
    if (flag) return 2;
    MPI_Init(&argc, &argv);
    MPI_Initialized(&flag);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    CHECK(flag == 1 && size == 4);

    // ring shift with Sendrecv
    int left = (rank + size - 1) % size, right = (rank + 1) % size, got = -1;
    MPI_Status st;
    MPI_Sendrecv(&rank, 1, MPI_INT, right, 3, &got, 1, MPI_INT, left, 3, MPI_COMM_WORLD, &st);
    CHECK(got == left && st.MPI_SOURCE == left && st.MPI_TAG == 3);

    // Irecv + Iprobe/Test polling + Waitany
    int a = -1, b = -1;
    MPI_Request reqs[2];
    MPI_Irecv(&a, 1, MPI_INT, left, 10, MPI_COMM_WORLD, &reqs[0]);
    MPI_Irecv(&b, 1, MPI_INT, right, 11, MPI_COMM_WORLD, &reqs[1]);
    int v10 = 100 + rank, v11 = 200 + rank;
    MPI_Send(&v10, 1, MPI_INT, right, 10, MPI_COMM_WORLD);
    MPI_Send(&v11, 1, MPI_INT, left, 11, MPI_COMM_WORLD);
    int done = 0;
    while (done < 2) {
        int idx = -5;
        MPI_Waitany(2, reqs, &idx, &st);
        CHECK(idx == 0 || idx == 1);
        ++done;
    }
    int idx = 0;
    MPI_Waitany(2, reqs, &idx, &st);
    CHECK(idx == MPI_UNDEFINED);
    CHECK(a == 100 + left && b == 200 + right);
    MPI_Request r;
    int c = -1;
    MPI_Irecv(&c, 1, MPI_INT, MPI_ANY_SOURCE, 12, MPI_COMM_WORLD, &r);
    MPI_Send(&rank, 1, MPI_INT, right, 12, MPI_COMM_WORLD);
    flag = 0;
    while (!flag) MPI_Test(&r, &flag, &st);
    CHECK(c == left && r == MPI_REQUEST_NULL);
    MPI_Barrier(MPI_COMM_WORLD);
    MPI_Iprobe(MPI_ANY_SOURCE, 99, MPI_COMM_WORLD, &flag, &st);
    CHECK(flag == 0);

    // Allgather, logical reductions
    std::vector<int> all(size, -1);
    int sq = rank * rank;
    MPI_Allgather(&sq, 1, MPI_INT, all.data(), 1, MPI_INT, MPI_COMM_WORLD);
    for (int i = 0; i < size; ++i) CHECK(all[i] == i * i);
    int yes = rank != 2, land = -1, lor = -1, bits = 1 << rank, bor = 0;
    MPI_Allreduce(&yes, &land, 1, MPI_INT, MPI_LAND, MPI_COMM_WORLD);
    MPI_Allreduce(&yes, &lor, 1, MPI_INT, MPI_LOR, MPI_COMM_WORLD);
    MPI_Allreduce(&bits, &bor, 1, MPI_INT, MPI_BOR, MPI_COMM_WORLD);
    CHECK(land == 0 && lor == 1 && bor == 15);

    // Comm_split / Comm_dup
    MPI_Comm half, dup, none;
    MPI_Comm_split(MPI_COMM_WORLD, rank % 2, -rank, &half);   // key reverses the order
    int hr = -1, hs = 0;
    MPI_Comm_rank(half, &hr);
    MPI_Comm_size(half, &hs);
    CHECK(hs == 2 && hr == (rank < 2 ? 1 : 0));
    MPI_Comm_split(MPI_COMM_WORLD, rank == 0 ? 7 : MPI_UNDEFINED, 0, &none);
    CHECK((rank == 0) == (none != MPI_COMM_NULL));
    MPI_Comm_dup(half, &dup);
    int hsum = 0;
    MPI_Allreduce(&rank, &hsum, 1, MPI_INT, MPI_SUM, dup);
    CHECK(hsum == (rank % 2 ? 4 : 2));
    MPI_Comm_free(&dup);
    MPI_Comm_free(&half);
    if (none != MPI_COMM_NULL) MPI_Comm_free(&none);

    // Dims_create (one dimension fixed), Cart_get, Cartdim_get
    int dims[2] = {0, 2}, periods[2] = {1, 0}, coords[2] = {-1, -1}, nd = 0;
    MPI_Dims_create(size, 2, dims);
    CHECK(dims[0] == 2 && dims[1] == 2);
    MPI_Comm cart;
    MPI_Cart_create(MPI_COMM_WORLD, 2, dims, periods, 0, &cart);
    int d2[2], p2[2];
    MPI_Cart_get(cart, 2, d2, p2, coords);
    MPI_Cartdim_get(cart, &nd);
    CHECK(nd == 2 && d2[0] == 2 && d2[1] == 2 && p2[0] == 1 && p2[1] == 0);
    CHECK(coords[0] == rank / 2 && coords[1] == rank % 2);
    MPI_Comm_free(&cart);

    // hvector + resized: every other column of a 4 x 6 double matrix, sent as 2 "columns"
    double m[4][6], colbuf[8];
    for (int y = 0; y < 4; ++y)
        for (int x = 0; x < 6; ++x) m[y][x] = 100 * rank + 10 * y + x;
    MPI_Datatype col, col_r;
    MPI_Aint a0, a1;
    MPI_Get_address(&m[0][0], &a0);
    MPI_Get_address(&m[1][0], &a1);
    MPI_Type_create_hvector(4, 1, a1 - a0, MPI_DOUBLE, &col);            // one column: 4 doubles, row pitch apart
    MPI_Type_create_resized(col, 0, 2 * (MPI_Aint)sizeof(double), &col_r);  // next element = 2 columns further
    MPI_Type_commit(&col_r);
    MPI_Sendrecv(&m[0][1], 2, col_r, right, 20, colbuf, 8, MPI_DOUBLE, left, 20, MPI_COMM_WORLD, &st);
    int cnt = 0;
    MPI_Get_count(&st, MPI_DOUBLE, &cnt);
    CHECK(cnt == 8);
    for (int k = 0; k < 2; ++k)
        for (int y = 0; y < 4; ++y) CHECK(colbuf[4 * k + y] == 100 * left + 10 * y + (1 + 2 * k));
    MPI_Type_free(&col_r);
    MPI_Type_free(&col);

    // Pack / Unpack / indexed_block: elements 1, 4, 7 of an int array travel as MPI_PACKED
    int src9[9], dst9[9], pos = 0, psize = 0, disp3[3] = {1, 4, 7};
    for (int i = 0; i < 9; ++i) { src9[i] = 10 * rank + i; dst9[i] = -1; }
    MPI_Datatype pick;
    MPI_Type_create_indexed_block(3, 1, disp3, MPI_INT, &pick);
    MPI_Type_commit(&pick);
    MPI_Pack_size(1, pick, MPI_COMM_WORLD, &psize);
    CHECK(psize == 12);
    char packed[64], inbox[64];
    double tagalong = 0.5 + rank, tag_in = 0;
    MPI_Pack(src9, 1, pick, packed, 64, &pos, MPI_COMM_WORLD);
    MPI_Pack(&tagalong, 1, MPI_DOUBLE, packed, 64, &pos, MPI_COMM_WORLD);
    CHECK(pos == 20);
    MPI_Sendrecv(packed, pos, MPI_PACKED, right, 30, inbox, 64, MPI_PACKED, left, 30, MPI_COMM_WORLD, &st);
    MPI_Get_count(&st, MPI_PACKED, &cnt);
    CHECK(cnt == 20);
    pos = 0;
    MPI_Unpack(inbox, cnt, &pos, dst9, 1, pick, MPI_COMM_WORLD);
    MPI_Unpack(inbox, cnt, &pos, &tag_in, 1, MPI_DOUBLE, MPI_COMM_WORLD);
    CHECK(dst9[1] == 10 * left + 1 && dst9[4] == 10 * left + 4 && dst9[7] == 10 * left + 7 && dst9[0] == -1);
    CHECK(tag_in == 0.5 + left && pos == 20);
    MPI_Type_free(&pick);

    int total = 0, v = 0, sv = 0;
    MPI_Get_version(&v, &sv);
    CHECK(v >= 1 && MPI_Wtick() > 0);
    MPI_Reduce(&fails, &total, 1, MPI_INT, MPI_SUM, 0, MPI_COMM_WORLD);
    if (rank == 0) std::printf("SHIM_EXTRA %s\n", total == 0 ? "PASSED" : "FAILED");
    MPI_Finalize();
    MPI_Finalized(&flag);
    return (total == 0 && flag) ? 0 : 1;
}
