/* Process-group plumbing: SelfComm, TcpComm, the global world(), mpi:: helpers and MpiTopology.
   Replaces the reference's direct MPI usage (include/stencil/mpi.hpp, mpi_topology.hpp; call sites SURVEY 2.4). */
#include "stencil/comm.hpp"
#include "stencil/mpi.hpp"
#include "stencil/mpi_topology.hpp"

#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <thread>

namespace stencil {

namespace {

/* tags >= kCollBase are reserved for the built-in collectives */
constexpr int kCollBase = 1 << 28;
constexpr int kTagGather = kCollBase + 1;
constexpr int kTagResult = kCollBase + 2;

[[noreturn]] void die(const char *what) {
  std::fprintf(stderr, "stencil::Comm fatal: %s (errno=%d %s)\n", what, errno, std::strerror(errno));
  std::exit(1);
}

double recv_timeout_s() {
  const char *e = std::getenv("STENCIL_COMM_TIMEOUT");
  return e ? std::atof(e) : 120.0;
}

/* (src, tag) -> FIFO of payloads */
class Mailbox {
  std::mutex mu_;
  std::map<std::pair<int, int>, std::deque<std::vector<unsigned char>>> q_;

public:
  void put(int src, int tag, std::vector<unsigned char> &&payload) {
    std::lock_guard<std::mutex> lk(mu_);
    q_[{src, tag}].push_back(std::move(payload));
  }
  bool take(int src, int tag, void *buf, size_t bytes) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = q_.find({src, tag});
    if (it == q_.end() || it->second.empty()) return false;
    std::vector<unsigned char> &m = it->second.front();
    if (m.size() != bytes) {
      std::fprintf(stderr, "stencil::Comm: size mismatch src=%d tag=%d got %zu want %zu\n", src, tag, m.size(), bytes);
      std::exit(1);
    }
    if (bytes) std::memcpy(buf, m.data(), bytes);
    it->second.pop_front();
    return true;
  }
};

class SelfComm : public Comm {
  Mailbox box_;

public:
  int rank() const override { return 0; }
  int size() const override { return 1; }
  const char *name() const override { return "self"; }
  void send(int dst, int tag, const void *buf, size_t bytes) override {
    if (dst != 0) die("SelfComm: send to non-zero rank");
    const unsigned char *p = static_cast<const unsigned char *>(buf);
    box_.put(0, tag, std::vector<unsigned char>(p, p + bytes));
  }
  bool try_recv(int src, int tag, void *buf, size_t bytes) override { return box_.take(src, tag, buf, bytes); }
  void barrier() override {}
  void allgather(const void *s, void *r, size_t n) override {
    if (n) std::memcpy(r, s, n);
  }
  void bcast(void *, size_t, int) override {}
  void allgatherv(const void *s, void *r, const std::vector<size_t> &c) override {
    if (c[0]) std::memcpy(r, s, c[0]);
  }
};

struct FrameHeader {
  int32_t dst, src, tag, pad;
  uint64_t bytes;
};

void write_all(int fd, const void *buf, size_t n) {
  const char *p = static_cast<const char *>(buf);
  while (n) {
    ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w < 0) {
      if (errno == EINTR) continue;
      die("socket write");
    }
    p += w;
    n -= size_t(w);
  }
}

/* returns false on orderly shutdown at a frame boundary */
bool read_all(int fd, void *buf, size_t n, bool eofOk) {
  char *p = static_cast<char *>(buf);
  size_t got = 0;
  while (got < n) {
    ssize_t r = ::recv(fd, p + got, n - got, 0);
    if (r == 0) {
      if (eofOk && got == 0) return false;
      die("socket closed mid-frame");
    }
    if (r < 0) {
      if (errno == EINTR) continue;
      if (eofOk) return false;
      die("socket read");
    }
    got += size_t(r);
  }
  return true;
}

/* Star topology: every non-root rank holds one socket to rank 0. Rank 0 relays frames whose dst != 0.
   One background thread per process drains sockets into the mailbox, so send() never deadlocks on a full pipe. */
class TcpComm : public Comm {
  int rank_, size_;
  std::vector<int> fds_;               // rank 0: fd per rank (fds_[0] unused); others: fds_[0] = to root
  std::vector<std::unique_ptr<std::mutex>> wmu_; // write lock per fd
  Mailbox box_;
  std::thread rx_;
  std::atomic<bool> stop_{false};
  int wake_[2] = {-1, -1};

  void tune(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  }

  void send_frame(int fd, std::mutex &mu, const FrameHeader &h, const void *payload) {
    std::lock_guard<std::mutex> lk(mu);
    write_all(fd, &h, sizeof(h));
    if (h.bytes) write_all(fd, payload, h.bytes);
  }

  void rx_loop() {
    std::vector<pollfd> pfds;
    std::vector<int> owner;
    if (rank_ == 0) {
      for (int r = 1; r < size_; ++r) {
        pfds.push_back({fds_[r], POLLIN, 0});
        owner.push_back(r);
      }
    } else {
      pfds.push_back({fds_[0], POLLIN, 0});
      owner.push_back(0);
    }
    pfds.push_back({wake_[0], POLLIN, 0});
    size_t live = pfds.size() - 1;
    while (!stop_.load() && live > 0) {
      int rc = ::poll(pfds.data(), pfds.size(), 1000);
      if (rc < 0) {
        if (errno == EINTR) continue;
        die("poll");
      }
      for (size_t i = 0; i + 1 < pfds.size(); ++i) {
        if (pfds[i].fd < 0 || !(pfds[i].revents & (POLLIN | POLLHUP | POLLERR))) continue;
        FrameHeader h;
        if (!read_all(pfds[i].fd, &h, sizeof(h), true)) {
          pfds[i].fd = -1; // peer went away
          --live;
          continue;
        }
        std::vector<unsigned char> payload(h.bytes);
        if (h.bytes) read_all(pfds[i].fd, payload.data(), h.bytes, false);
        if (h.dst == rank_) {
          box_.put(h.src, h.tag, std::move(payload));
        } else if (rank_ == 0 && h.dst > 0 && h.dst < size_) {
          send_frame(fds_[h.dst], *wmu_[h.dst], h, payload.data());
        } else {
          die("frame for unknown destination");
        }
      }
    }
  }

public:
  TcpComm(int rank, int size, const std::string &addr, int port) : rank_(rank), size_(size) {
    if (::pipe(wake_) != 0) die("pipe");
    if (rank_ == 0) {
      fds_.assign(size_, -1);
      int ls = ::socket(AF_INET, SOCK_STREAM, 0);
      if (ls < 0) die("socket");
      int one = 1;
      setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
      sockaddr_in sa{};
      sa.sin_family = AF_INET;
      sa.sin_addr.s_addr = htonl(INADDR_ANY);
      sa.sin_port = htons(uint16_t(port));
      if (::bind(ls, reinterpret_cast<sockaddr *>(&sa), sizeof(sa)) != 0) die("bind");
      if (::listen(ls, size_) != 0) die("listen");
      for (int joined = 1; joined < size_; ++joined) {
        int fd = ::accept(ls, nullptr, nullptr);
        if (fd < 0) die("accept");
        tune(fd);
        int32_t who = -1;
        read_all(fd, &who, sizeof(who), false);
        if (who <= 0 || who >= size_ || fds_[who] != -1) die("bad rank in hello");
        fds_[who] = fd;
      }
      ::close(ls);
    } else {
      fds_.assign(1, -1);
      addrinfo hints{}, *res = nullptr;
      hints.ai_family = AF_INET;
      hints.ai_socktype = SOCK_STREAM;
      const std::string ports = std::to_string(port);
      if (getaddrinfo(addr.c_str(), ports.c_str(), &hints, &res) != 0 || !res) die("getaddrinfo");
      const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(120);
      int fd = -1;
      for (;;) {
        fd = ::socket(AF_INET, SOCK_STREAM, 0);
        if (fd < 0) die("socket");
        if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
        ::close(fd);
        if (std::chrono::steady_clock::now() > deadline) die("connect to rank 0 timed out");
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
      }
      freeaddrinfo(res);
      tune(fd);
      int32_t who = rank_;
      write_all(fd, &who, sizeof(who));
      fds_[0] = fd;
    }
    for (size_t i = 0; i < fds_.size(); ++i) wmu_.emplace_back(new std::mutex);
    rx_ = std::thread([this] { rx_loop(); });
    barrier();
  }

  ~TcpComm() override {
    barrier();
    stop_.store(true);
    char c = 0;
    (void)!::write(wake_[1], &c, 1);
    for (int fd : fds_)
      if (fd >= 0) ::shutdown(fd, SHUT_RDWR);
    if (rx_.joinable()) rx_.join();
    for (int fd : fds_)
      if (fd >= 0) ::close(fd);
    ::close(wake_[0]);
    ::close(wake_[1]);
  }

  int rank() const override { return rank_; }
  int size() const override { return size_; }
  const char *name() const override { return "tcp"; }

  void send(int dst, int tag, const void *buf, size_t bytes) override {
    if (dst == rank_) {
      const unsigned char *p = static_cast<const unsigned char *>(buf);
      box_.put(rank_, tag, std::vector<unsigned char>(p, p + bytes));
      return;
    }
    FrameHeader h{dst, rank_, tag, 0, bytes};
    if (rank_ == 0)
      send_frame(fds_[dst], *wmu_[dst], h, buf);
    else
      send_frame(fds_[0], *wmu_[0], h, buf);
  }
  bool try_recv(int src, int tag, void *buf, size_t bytes) override { return box_.take(src, tag, buf, bytes); }
};

std::unique_ptr<Comm> g_world;
std::mutex g_worldMu;

} // namespace

/* ---- default collectives over p2p through rank 0 ---- */

void Comm::recv(int src, int tag, void *buf, size_t bytes) {
  const double limit = recv_timeout_s();
  const double t0 = wtime();
  unsigned spins = 0;
  while (!try_recv(src, tag, buf, bytes)) {
    if (++spins > 64) std::this_thread::sleep_for(std::chrono::microseconds(20));
    if (wtime() - t0 > limit) {
      std::fprintf(stderr, "stencil::Comm rank %d: recv(src=%d, tag=%d, %zu B) timed out after %.0f s\n", rank(), src, tag,
                   bytes, limit);
      std::exit(1);
    }
  }
}

void Comm::allgatherv(const void *sendbuf, void *recvbuf, const std::vector<size_t> &counts) {
  const int n = size(), me = rank();
  size_t total = 0;
  std::vector<size_t> off(n);
  for (int r = 0; r < n; ++r) {
    off[r] = total;
    total += counts[r];
  }
  char *out = static_cast<char *>(recvbuf);
  if (me == 0) {
    if (counts[0]) std::memcpy(out, sendbuf, counts[0]);
    for (int r = 1; r < n; ++r) recv(r, kTagGather, out + off[r], counts[r]);
    for (int r = 1; r < n; ++r) send(r, kTagResult, out, total);
  } else {
    send(0, kTagGather, sendbuf, counts[me]);
    recv(0, kTagResult, out, total);
  }
}

void Comm::allgather(const void *sendbuf, void *recvbuf, size_t bytesPerRank) {
  allgatherv(sendbuf, recvbuf, std::vector<size_t>(size_t(size()), bytesPerRank));
}

void Comm::barrier() {
  char token = 0;
  std::vector<char> all(static_cast<size_t>(size()));
  allgather(&token, all.data(), 1);
}

void Comm::bcast(void *buf, size_t bytes, int root) {
  /* small set-up payloads only: gather-everything keeps the code to one primitive */
  std::vector<char> all(bytes * size_t(size()));
  allgather(buf, all.data(), bytes);
  if (rank() != root && bytes) std::memcpy(buf, all.data() + size_t(root) * bytes, bytes);
}

double Comm::allreduce_max(double v) {
  std::vector<double> all(static_cast<size_t>(size()));
  allgather(&v, all.data(), sizeof(double));
  double m = all[0];
  for (double e : all) m = e > m ? e : m;
  return m;
}

uint64_t Comm::allreduce_sum(uint64_t v) {
  std::vector<uint64_t> all(static_cast<size_t>(size()));
  allgather(&v, all.data(), sizeof(uint64_t));
  uint64_t s = 0;
  for (uint64_t e : all) s += e;
  return s;
}

std::string Comm::hostname() const {
  const char *fake = std::getenv("STENCIL_FAKE_HOSTNAME"); // lets one box pose as several nodes in tests
  if (fake) return fake;
  char buf[256] = {0};
  if (gethostname(buf, sizeof(buf) - 1) != 0) return "localhost";
  return buf;
}

double Comm::wtime() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

std::unique_ptr<Comm> make_self_comm() { return std::unique_ptr<Comm>(new SelfComm); }

std::unique_ptr<Comm> make_tcp_comm(int rank, int size, const std::string &addr, int port) {
  return std::unique_ptr<Comm>(new TcpComm(rank, size, addr, port));
}

Comm &world() {
  std::lock_guard<std::mutex> lk(g_worldMu);
  if (!g_world) {
    const char *ws = std::getenv("WORLD_SIZE");
    const char *rk = std::getenv("RANK");
    const char *ma = std::getenv("MASTER_ADDR");
    const char *off = std::getenv("STENCIL_NO_ENV_COMM");
    /* under an MPI launcher (or when asked) and built with USE_MPI: MPI; else torchrun-style env: TCP; else alone */
    const char *want = std::getenv("STENCIL_COMM");
    const bool mpiLaunch = std::getenv("OMPI_COMM_WORLD_SIZE") || std::getenv("PMI_SIZE") || std::getenv("PMIX_RANK") || std::getenv("SLURM_PROCID");
    if (!off && ((want && std::string(want) == "mpi") || (!want && mpiLaunch))) g_world = make_mpi_comm(); // null without USE_MPI
    if (g_world) {
    } else if (!off && ws && rk && ma && std::atoi(ws) > 1) {
      int port = 29517;
      if (const char *sp = std::getenv("STENCIL_PORT"))
        port = std::atoi(sp);
      else if (const char *mp = std::getenv("MASTER_PORT"))
        port = std::atoi(mp) + 17;
      g_world = make_tcp_comm(std::atoi(rk), std::atoi(ws), ma, port);
    } else {
      g_world = make_self_comm();
    }
  }
  return *g_world;
}

void set_world(std::unique_ptr<Comm> c) {
  std::lock_guard<std::mutex> lk(g_worldMu);
  g_world = std::move(c);
}

void finalize_world() {
  std::lock_guard<std::mutex> lk(g_worldMu);
  g_world.reset();
}

} // namespace stencil

namespace mpi {
int world_rank() { return stencil::world().rank(); }
int world_size() { return stencil::world().size(); }
} // namespace mpi

/* ---- MpiTopology ---- */

MpiTopology::MpiTopology(stencil::Comm &comm) : comm_(&comm), me_(comm.rank()) {
  constexpr size_t kName = 256;
  char mine[kName] = {0};
  std::strncpy(mine, comm.hostname().c_str(), kName - 1);
  std::vector<char> all(kName * size_t(comm.size()));
  comm.allgather(mine, all.data(), kName);
  for (int r = 0; r < comm.size(); ++r) hosts_.emplace_back(all.data() + size_t(r) * kName);
  build();
}

MpiTopology::MpiTopology(const std::vector<std::string> &hosts, int me) : comm_(nullptr), hosts_(hosts), me_(me) { build(); }

void MpiTopology::build() {
  std::map<std::string, int> number;
  nodeOfRank_.resize(hosts_.size());
  for (size_t r = 0; r < hosts_.size(); ++r) {
    auto ins = number.emplace(hosts_[r], int(number.size()));
    nodeOfRank_[r] = ins.first->second;
  }
  numNodes_ = int(number.size());
  for (size_t r = 0; r < hosts_.size(); ++r)
    if (nodeOfRank_[r] == nodeOfRank_[size_t(me_)]) colocated_.push_back(int(r));
}

int MpiTopology::colocated_rank() const noexcept {
  for (size_t i = 0; i < colocated_.size(); ++i)
    if (colocated_[i] == me_) return int(i);
  return 0;
}

bool MpiTopology::colocated(int rank) const noexcept {
  return rank >= 0 && size_t(rank) < nodeOfRank_.size() && nodeOfRank_[size_t(rank)] == nodeOfRank_[size_t(me_)];
}
