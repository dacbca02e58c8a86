// Datatypes, groups and topologies against the C++ API of b200comm, ranks = threads.
//   mpi7.cpp   indexed type (4 floats @5, 2 floats @12) sent to every rank incl. the
//              root itself, received as 6 contiguous floats
//   mpi8.cpp   struct Particle {4 float, 2 int}: particle i goes to rank i
//   mpi9.cpp   split the world in halves via groups, Allreduce per half and on world
//   mpi10.cpp  non-periodic sqrt(N) x sqrt(N) Cartesian grid, 4-neighbour exchange
//   mpi-complex-types.cpp  hindexed over a 1-D subarray across three separate buffers
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "b200comm/comm.hpp"

using namespace b200;
using namespace b200::comm;

struct Particle { float x, y, z, velocity; int id, type; };

static void rank_main(Comm world, std::vector<std::string>* out) {
    const int me = world.rank(), n = world.size();
    std::ostringstream os;
    os << "rank " << me << ":\t";
    const Datatype FLOAT = Datatype::basic(Basic::FLOAT), INT = Datatype::basic(Basic::INT32);

    // ---- mpi7
    Datatype idx = Datatype::indexed({4, 2}, {5, 12}, FLOAT);
    idx.commit();
    float a[16];
    for (int i = 0; i < 16; ++i) a[i] = (float)i;
    std::vector<Request> sends;
    if (me == 0)
        for (int r = 0; r < n; ++r) sends.push_back(world.isend(a, 1, idx, r, 1));
    float b[6] = {0};
    world.recv(b, 6, FLOAT, 0, 1);
    waitall(sends);
    for (float v : b) os << v << ",";

    // ---- mpi8
    Datatype pt = Datatype::create_struct({4, 2}, {0, (int64_t)offsetof(Particle, id)}, {FLOAT, INT});
    pt.commit();
    std::vector<Particle> ps((size_t)n);
    sends.clear();
    if (me == 0)
        for (int r = 0; r < n; ++r) {
            ps[(size_t)r] = Particle{1.f * r, -1.f * r, 2.f * r, 0.25f, r, r % 2};
            sends.push_back(world.isend(&ps[(size_t)r], 1, pt, r, 2));
        }
    Particle p{};
    world.recv(&p, 1, pt, 0, 2);
    waitall(sends);
    os << " particle(" << p.x << "," << p.y << "," << p.z << "," << p.velocity << "," << p.id << "," << p.type << ")";

    // ---- mpi9
    const int half = n / 2;
    std::vector<int> lo, hi;
    for (int r = 0; r < n; ++r) (r < half ? lo : hi).push_back(r);
    Comm c_lo = world.create(world.group().incl(lo)), c_hi = world.create(world.group().incl(hi));
    Comm mine = me < half ? c_lo : c_hi;
    int s_half = 0, s_world = 0;
    mine.allreduce(&me, &s_half, 1, INT, Op::SUM);
    world.allreduce(&me, &s_world, 1, INT, Op::SUM);
    os << " half-sum=" << s_half << " world-sum=" << s_world;

    // ---- mpi10
    const int dim = (int)std::lround(std::sqrt((double)n));
    if (dim * dim == n) {
        Comm cart = world.cart_create({dim, dim}, {0, 0});
        int up, down, left, right;
        cart.cart_shift(0, 1, &up, &down);
        cart.cart_shift(1, 1, &left, &right);
        int nb[4] = {up, down, left, right}, got[4] = {-1, -1, -1, -1};
        const int id = cart.rank();
        std::vector<Request> rq;
        for (int k = 0; k < 4; ++k) {
            rq.push_back(cart.irecv(&got[k], 1, INT, nb[k], 3));
            rq.push_back(cart.isend(&id, 1, INT, nb[k], 3));
        }
        waitall(rq);
        std::vector<int> co = cart.cart_coords(id);
        os << " cart(" << co[0] << "," << co[1] << ") nbrs " << got[0] << " " << got[1] << " " << got[2] << " " << got[3];
    }

    // ---- mpi-complex-types (ranks 0 -> 1)
    if (n >= 2 && me < 2) {
        int *B1 = new int[1500], *B2 = new int[8], *B3 = new int[28];
        for (int i = 0; i < 8; ++i) {
            B1[i] = me == 0 ? i : -1; B2[i] = me == 0 ? 2 * i : -1; B3[i] = me == 0 ? 2 * i + 1 : -1;
        }
        Datatype sub = Datatype::subarray({8}, {3}, {me == 0 ? 3 : 0}, Order::C, INT);
        sub.commit();
        Datatype fin = Datatype::hindexed({1, 1, 1},
                                          {0, (int64_t)((char*)B2 - (char*)B1), (int64_t)((char*)B3 - (char*)B1)}, sub);
        fin.commit();
        if (me == 0) world.send(B1, 1, fin, 1, 123);
        else {
            world.recv(B1, 1, fin, 0, 123);
            os << " complex:";
            for (int* B : {B1, B2, B3}) { os << " ["; for (int i = 0; i < 4; ++i) os << B[i] << (i < 3 ? " " : ""); os << "]"; }
        }
        delete[] B1; delete[] B2; delete[] B3;
    }
    (*out)[(size_t)me] = os.str();
    finalize(world);
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 4;
    std::vector<Comm> comms = init_local(n);
    std::vector<std::string> out((size_t)n);
    std::vector<std::thread> th;
    for (int r = 0; r < n; ++r) th.emplace_back(rank_main, comms[(size_t)r], &out);
    for (auto& t : th) t.join();
    for (auto& s : out) std::printf("%s\n", s.c_str());
    return 0;
}
