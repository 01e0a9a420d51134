#include "transport.h"
#include "threads.h"

#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <chrono>
#include <cstring>

#include "uda_rt.h"

namespace uda {

class Connection {
public:
    explicit Connection(int fd) : fd_(fd) {
        int one = 1;
        setsockopt(fd_, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    }
    ~Connection() { close_fd(); }
    bool send_frame(uint8_t type, uint8_t credits, uint64_t src_req, const std::string& text, const uint8_t* payload, size_t plen) {
        MsgHeader h{};
        h.type = type;
        h.credits = credits;
        h.tot_len = (uint32_t)text.size();
        h.src_req = src_req;
        h.payload_len = plen;
        std::lock_guard<std::mutex> lk(wmu_);
        return write_all(&h, sizeof(h)) && write_all(text.data(), text.size()) && (plen == 0 || write_all(payload, plen));
    }
    bool read_all(void* buf, size_t n) {
        uint8_t* p = (uint8_t*)buf;
        while (n) {
            ssize_t r = ::recv(fd_, p, n, 0);
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) return false;
            p += r;
            n -= (size_t)r;
        }
        return true;
    }
    bool discard(size_t n) {
        uint8_t tmp[4096];
        while (n) {
            size_t k = n < sizeof(tmp) ? n : sizeof(tmp);
            if (!read_all(tmp, k)) return false;
            n -= k;
        }
        return true;
    }
    void shutdown_rw() {
        if (fd_ >= 0) ::shutdown(fd_, SHUT_RDWR);
    }
    void close_fd() {
        if (fd_ >= 0) { ::close(fd_); fd_ = -1; }
    }

private:
    bool write_all(const void* buf, size_t n) {
        const uint8_t* p = (const uint8_t*)buf;
        while (n) {
            ssize_t r = ::send(fd_, p, n, MSG_NOSIGNAL);
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) return false;
            p += r;
            n -= (size_t)r;
        }
        return true;
    }
    int fd_;
    std::mutex wmu_;
};

// ------------------------------------------------------------------------------------------ server
ShuffleServer::ShuffleServer(int port, RequestHandler h) : port_(port), on_request_(std::move(h)) {}
ShuffleServer::~ShuffleServer() { stop(); }

void ShuffleServer::start() {
    listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    if (listen_fd_ < 0) throw UdaError("socket() failed");
    int one = 1;
    setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in sa{};
    sa.sin_family = AF_INET;
    sa.sin_addr.s_addr = htonl(INADDR_ANY);
    sa.sin_port = htons((uint16_t)port_);
    if (::bind(listen_fd_, (sockaddr*)&sa, sizeof(sa)) != 0 || ::listen(listen_fd_, 128) != 0) {
        ::close(listen_fd_);
        listen_fd_ = -1;
        UDA_LOG(lsERROR, "cannot listen on port %d: %s", port_, strerror(errno));
        throw UdaError("cannot bind/listen on the data port");
    }
    socklen_t sl = sizeof(sa);
    if (getsockname(listen_fd_, (sockaddr*)&sa, &sl) == 0) port_ = ntohs(sa.sin_port);
    running_ = true;
    accept_thread_ = spawn_guarded("accept", nullptr, nullptr, [this] { accept_loop(); });
    UDA_LOG(lsINFO, "shuffle server listening on port %d", port_);
}

void ShuffleServer::accept_loop() {
    while (running_) {
        int fd = ::accept(listen_fd_, nullptr, nullptr);
        if (fd < 0) {
            if (errno == EINTR) continue;
            break;  // listen socket closed by stop()
        }
        auto c = std::make_shared<Connection>(fd);
        std::lock_guard<std::mutex> lk(mu_);
        conns_.push_back(c);
        conn_threads_.emplace_back([this, c] { serve(c); });
    }
}

void ShuffleServer::serve(std::shared_ptr<Connection> c) {
    MsgHeader h;
    std::string text;
    while (running_ && c->read_all(&h, sizeof(h))) {
        if (h.tot_len > kMaxMsgText) { UDA_LOG(lsERROR, "oversized message (%u bytes): dropping connection", h.tot_len); break; }
        text.resize(h.tot_len);
        if (h.tot_len && !c->read_all(&text[0], h.tot_len)) break;
        if (h.payload_len && !c->discard(h.payload_len)) break;
        if (h.type == MSG_NOOP) continue;  // pure credit return
        if (h.type != MSG_RTS) { UDA_LOG(lsWARN, "unexpected message type 0x%x", h.type); continue; }
        FetchRequest req;
        if (!parse_fetch_request(text, &req)) { UDA_LOG(lsERROR, "malformed fetch request '%s'", text.c_str()); continue; }
        on_request_(req, c);
    }
    c->shutdown_rw();  // marks the connection bad: later replies fail and the chunk is released by the caller
}

bool ShuffleServer::reply(const std::shared_ptr<Connection>& c, uint64_t src_req, const std::string& ack, const uint8_t* payload, size_t plen) {
    return c->send_frame(MSG_RTS, 1, src_req, ack, payload, plen);
}

void ShuffleServer::stop() {
    if (!running_.exchange(false)) return;
    // wake accept() with shutdown(), join, and only then close: closing first would let the descriptor number be
    // reused while the accept thread may still be about to call accept() on it
    const int lfd = listen_fd_.load();
    if (lfd >= 0) ::shutdown(lfd, SHUT_RDWR);
    if (accept_thread_.joinable()) accept_thread_.join();
    if (lfd >= 0) { ::close(lfd); listen_fd_ = -1; }
    std::vector<std::thread> ts;
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto& c : conns_) c->shutdown_rw();
        ts.swap(conn_threads_);
    }
    for (auto& t : ts) t.join();
    std::lock_guard<std::mutex> lk(mu_);
    conns_.clear();
}

// ------------------------------------------------------------------------------------------ client
TcpInputClient::TcpInputClient(int port, int wqes) : port_(port), wqes_(wqes < 2 ? 2 : wqes) {}
TcpInputClient::~TcpInputClient() { stop(); }

static int connect_with_retries(const std::string& host, int port) {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) != 0 || !res) return -1;
    int fd = -1;
    for (int attempt = 0; attempt < 5 && fd < 0; ++attempt) {  // RDMA_DEFAULT_RNR_RETRY-style bounded reconnects
        fd = ::socket(AF_INET, SOCK_STREAM, 0);
        if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) != 0) {
            ::close(fd);
            fd = -1;
            std::this_thread::sleep_for(std::chrono::milliseconds(50 * (attempt + 1)));
        }
    }
    freeaddrinfo(res);
    return fd;
}

std::shared_ptr<TcpInputClient::Peer> TcpInputClient::peer_for(const std::string& host) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = peers_.find(host);
    if (it != peers_.end() && !it->second->bad) return it->second;
    int fd = connect_with_retries(host, port_);
    if (fd < 0) {
        UDA_LOG(lsERROR, "could not connect to host %s on port %d", host.c_str(), port_);
        return nullptr;
    }
    auto p = std::make_shared<Peer>();
    p->conn = std::make_shared<Connection>(fd);
    p->credits = wqes_ - 1;
    p->reader = spawn_guarded("client-read", nullptr, nullptr, [this, p] { read_loop(p); });
    if (it != peers_.end()) { if (it->second->reader.joinable()) it->second->reader.detach(); }
    peers_[host] = p;
    return p;
}

void TcpInputClient::fetch(const std::string& host, FetchRequest req, uint8_t* dest, size_t cap, Done done) {
    if (req.chunk_size <= 0) {
        UDA_LOG(lsERROR, "illegal fetch request size of %d bytes", req.chunk_size);  // regression greps this line
        throw UdaError("illegal fetch request size of 0 or less bytes");
    }
    auto p = peer_for(host);
    if (!p) throw UdaError("could not connect to the map-output provider");
    uint64_t cookie = next_cookie_++;
    req.request_ptr = cookie;
    req.remote_addr = (uint64_t)(uintptr_t)dest;
    std::string text = format_fetch_request(req);
    bool send_now = false, failed = false;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        if (p->bad) failed = true;
        else {
            p->inflight[cookie] = Pending{dest, cap, std::move(done)};
            if (p->credits > 0) { --p->credits; send_now = true; }
            else p->backlog.emplace_back(text, cookie);  // no credit: wait for an ack to free one
        }
    }
    if (failed) throw UdaError("connection to the map-output provider is down");
    if (send_now && !p->conn->send_frame(MSG_RTS, 0, 0, text, nullptr, 0)) fail_all(p);
}

void TcpInputClient::read_loop(std::shared_ptr<Peer> p) {
    MsgHeader h;
    std::string text;
    while (!stopping_ && p->conn->read_all(&h, sizeof(h))) {
        if (h.tot_len > kMaxMsgText) break;
        text.resize(h.tot_len);
        if (h.tot_len && !p->conn->read_all(&text[0], h.tot_len)) break;
        Pending pend{};
        bool known = false;
        std::string next_text;
        {
            std::lock_guard<std::mutex> lk(p->mu);
            auto it = p->inflight.find(h.src_req);
            if (it != p->inflight.end()) { pend = std::move(it->second); p->inflight.erase(it); known = true; }
            // the ack returns a credit: use it for the oldest backlogged request, else bank it
            if (!p->backlog.empty()) { next_text = p->backlog.front().first; p->backlog.erase(p->backlog.begin()); }
            else ++p->credits;
        }
        bool ok = true;
        if (h.payload_len) {
            if (known && h.payload_len <= pend.cap) ok = p->conn->read_all(pend.dest, h.payload_len);
            else { ok = p->conn->discard(h.payload_len) && !known; }
        }
        if (!next_text.empty()) p->conn->send_frame(MSG_RTS, 1, 0, next_text, nullptr, 0);
        if (known) {
            FetchAck ack;
            bool parsed = parse_fetch_ack(text, &ack);
            // the ack text comes off a socket too: what it claims to have sent must be what actually followed it
            if (parsed && ok && ack.sent != (long long)h.payload_len) {
                UDA_LOG(lsERROR, "fetch ack claims %lld bytes, the frame carried %llu", ack.sent, (unsigned long long)h.payload_len);
                parsed = false;
            }
            pend.done(ok && parsed, ack);
        }
        if (!ok) break;
    }
    fail_all(p);
}

void TcpInputClient::fail_all(const std::shared_ptr<Peer>& p) {
    std::map<uint64_t, Pending> victims;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->bad = true;
        victims.swap(p->inflight);
        p->backlog.clear();
    }
    p->conn->shutdown_rw();
    for (auto& kv : victims) kv.second.done(false, FetchAck{});
}

void TcpInputClient::stop() {
    if (stopping_.exchange(true)) return;
    std::map<std::string, std::shared_ptr<Peer>> ps;
    {
        std::lock_guard<std::mutex> lk(mu_);
        ps.swap(peers_);
    }
    for (auto& kv : ps) {
        kv.second->conn->shutdown_rw();
        if (kv.second->reader.joinable()) kv.second->reader.join();
    }
}

// ---- fetcher factory ------------------------------------------------------------------------------------
namespace {
std::mutex g_factory_mu;
InputClientFactory g_factory;
}  // namespace

void set_input_client_factory(InputClientFactory f) {
    std::lock_guard<std::mutex> lk(g_factory_mu);
    g_factory = std::move(f);
}

std::unique_ptr<InputClient> make_input_client(int port, int wqes_per_conn) {
    InputClientFactory f;
    {
        std::lock_guard<std::mutex> lk(g_factory_mu);
        f = g_factory;
    }
    if (f) return f(port, wqes_per_conn);
    return std::make_unique<TcpInputClient>(port, wqes_per_conn);
}

}  // namespace uda
