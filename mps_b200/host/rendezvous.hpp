// rendezvous.hpp — one process per GPU: how the ranks of a multi-GPU run find each other from C++.
// The launcher (torchrun, mpirun, srun or a shell loop) sets RANK, WORLD_SIZE, LOCAL_RANK, MASTER_ADDR and
// MASTER_PORT; rank 0 creates the 128-byte NCCL unique id (mpsb_comm_unique_id) and hands it to every other
// rank over a plain TCP connection on MASTER_PORT + 1 (MASTER_PORT itself is left to the launcher's own
// rendezvous). That id is all mpsb_attach_comm needs; every later exchange is device-side.
#pragma once

#include <arpa/inet.h>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <netinet/in.h>
#include <string>
#include <sys/socket.h>
#include <thread>
#include <unistd.h>

namespace rendezvous {

struct Env {
    int rank = 0, world = 1, localRank = 0, port = 29500;
    std::string addr = "127.0.0.1";
};

inline Env fromEnvironment() {
    Env e;
    if (const char* v = std::getenv("RANK")) e.rank = std::atoi(v);
    if (const char* v = std::getenv("WORLD_SIZE")) e.world = std::atoi(v);
    e.localRank = e.rank;
    if (const char* v = std::getenv("LOCAL_RANK")) e.localRank = std::atoi(v);
    if (const char* v = std::getenv("MASTER_ADDR")) e.addr = v;
    if (const char* v = std::getenv("MASTER_PORT")) e.port = std::atoi(v);
    if (e.world < 1) e.world = 1;
    return e;
}

inline bool sendAll(int fd, const unsigned char* p, size_t n) {
    while (n > 0) {
        ssize_t k = ::send(fd, p, n, 0);
        if (k <= 0) return false;
        p += k;
        n -= (size_t) k;
    }
    return true;
}
inline bool recvAll(int fd, unsigned char* p, size_t n) {
    while (n > 0) {
        ssize_t k = ::recv(fd, p, n, 0);
        if (k <= 0) return false;
        p += k;
        n -= (size_t) k;
    }
    return true;
}

// Rank 0 passes the id it made; the others receive it into `id`. False on any socket failure or after 120 s.
inline bool shareUniqueId(const Env& e, unsigned char id[128]) {
    if (e.world == 1) return true;
    const int port = e.port + 1;
    if (e.rank == 0) {
        int srv = ::socket(AF_INET, SOCK_STREAM, 0);
        if (srv < 0) return false;
        int one = 1;
        ::setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        sockaddr_in a{};
        a.sin_family      = AF_INET;
        a.sin_addr.s_addr = htonl(INADDR_ANY);
        a.sin_port        = htons((unsigned short) port);
        if (::bind(srv, (sockaddr*) &a, sizeof a) != 0 || ::listen(srv, e.world) != 0) {
            ::close(srv);
            return false;
        }
        bool ok = true;
        for (int k = 1; k < e.world && ok; k++) {
            int c = ::accept(srv, nullptr, nullptr);
            ok    = c >= 0 && sendAll(c, id, 128);
            if (c >= 0) ::close(c);
        }
        ::close(srv);
        return ok;
    }
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(120);
    while (std::chrono::steady_clock::now() < deadline) {
        int c = ::socket(AF_INET, SOCK_STREAM, 0);
        if (c < 0) return false;
        sockaddr_in a{};
        a.sin_family = AF_INET;
        a.sin_port   = htons((unsigned short) port);
        if (::inet_pton(AF_INET, e.addr.c_str(), &a.sin_addr) != 1) {
            ::close(c);
            return false;
        }
        if (::connect(c, (sockaddr*) &a, sizeof a) == 0) {
            bool ok = recvAll(c, id, 128);
            ::close(c);
            return ok;
        }
        ::close(c);
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    return false;
}

} // namespace rendezvous
