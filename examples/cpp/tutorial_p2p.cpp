// Point-to-point tutorial against the C++ API of b200comm, ranks = threads.
//
// Covers the behaviours of the reference's first six tutorials with ONE program:
//   mpi1.cpp  init / rank / size / processor name / finalize
//   mpi2.cpp  error policy (exceptions with message + error class)
//   mpi3.cpp  blocking send/recv, receiver sizes its buffer with probe + get_count
//   mpi4.cpp  ping-pong counter up to 10
//   mpi5.cpp  non-blocking neighbour exchange on an open chain
//   mpi6.cpp  ... followed by a gather of (self, prev, next) to the root
// Build:  g++ -std=c++17 -I include examples/cpp/tutorial_p2p.cpp \
//             -L cuda_mpi_scratch_b200/lib -lb200comm -Wl,-rpath,$PWD/cuda_mpi_scratch_b200/lib -pthread
// Each rank prints ONE string (the reference's idiom against interleaving, mpi7.cpp:56-62).
#include <cstdio>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "b200comm/comm.hpp"

using namespace b200;
using namespace b200::comm;

static void rank_main(Comm world, std::vector<std::string>* out) {
    const int me = world.rank(), n = world.size();
    std::ostringstream os;
    os << processor_name() << " - rank " << me << " of " << n << ": ";
    const Datatype CHAR = Datatype::basic(Basic::CHAR), INT = Datatype::basic(Basic::INT32);

    // ---- mpi3: strings between rank 0 and 1, size discovered by probe
    if (me < 2) {
        const int peer = 1 - me;
        const std::string msg = "Hello from rank " + std::to_string(me);
        auto recv_string = [&]() {
            Status st = world.probe(peer, 7);
            std::vector<char> buf((size_t)st.count(CHAR));
            world.recv(buf.data(), (int64_t)buf.size(), CHAR, peer, 7);
            return std::string(buf.begin(), buf.end());
        };
        std::string got;
        if (me == 0) { world.send(msg.data(), (int64_t)msg.size(), CHAR, peer, 7); got = recv_string(); }
        else { got = recv_string(); world.send(msg.data(), (int64_t)msg.size(), CHAR, peer, 7); }
        os << "[" << got << "] ";
        // ---- mpi4: bounce a counter until it reaches 10
        int v = 0, turn = 0;
        while (v < 10) {
            if (turn % 2 == me) { ++v; world.send(&v, 1, INT, peer, 1); }
            else world.recv(&v, 1, INT, peer, 1);
            ++turn;
        }
        os << "counter=" << v << " ";
    }
    // ---- mpi5: open chain, Isend/Irecv to rank +-1, edges are PROC_NULL
    int prev = -1, next = -1;
    const int left = me > 0 ? me - 1 : PROC_NULL, right = me < n - 1 ? me + 1 : PROC_NULL;
    std::vector<Request> reqs;
    reqs.push_back(world.irecv(&prev, 1, INT, left, 11));
    reqs.push_back(world.irecv(&next, 1, INT, right, 12));
    reqs.push_back(world.isend(&me, 1, INT, left, 12));
    reqs.push_back(world.isend(&me, 1, INT, right, 11));
    waitall(reqs);
    os << "prev=" << prev << " next=" << next << " ";
    // ---- mpi6: gather the triples on the root (non-roots pass no receive buffer)
    int triple[3] = {me, prev, next};
    std::vector<int> all(me == 0 ? 3 * (size_t)n : 0);
    world.gather(triple, 3, INT, me == 0 ? all.data() : nullptr, 3, INT, 0);
    if (me == 0) {
        os << "gathered:";
        for (int r = 0; r < n; ++r) os << " (" << all[3 * r] << "," << all[3 * r + 1] << "," << all[3 * r + 2] << ")";
    }
    // ---- mpi2: an error is an exception carrying message and error class
    try {
        world.send(&me, 1, INT, n + 5, 0);
    } catch (const CommError& e) {
        os << " [caught error class " << e.code() << ": " << error_class_string(e.code()) << "]";
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
