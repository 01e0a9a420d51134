// CPU plumbing transport: the fetch-request / data+ack exchange of the reference's DataNet layer over
// TCP (loop-back or LAN), used for BASELINE config 1 and for API-compatible deployments without
// NVLink peers. The GPU data path does not go through here (see csrc/kernels/shuffle.cu).
//
// Reference behaviour covered (src/DataNet/):
//   * message header {type, credits, tot_len, src_req} + <= 800 bytes of text  (RDMAComm.h:41-47,65-72)
//   * call site A: client sends the fetch-request text                        (RDMAClient.cc:559-600)
//   * call site B: server sends the chunk, then the ack text echoing src_req   (RDMAServer.cc:537-631);
//     on TCP the chunk bytes follow the ack in the same frame (payload_len), where the reference
//     RDMA-WRITEs them first
//   * call site C: credits — at most (wqes-1) un-acked requests per connection, the rest wait in a
//     backlog that is drained as acks return                                  (RDMAClient.cc:43-125)
//   * one connection per remote host, cached; up to 5 connect attempts        (RDMAClient.cc:215-356)
//   * a bad connection fails its in-flight requests instead of hanging        (RDMAServer.cc:173-196)
#pragma once
#include <atomic>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "protocol.h"

namespace uda {

enum MsgType : uint8_t { MSG_NOOP = 0x01, MSG_RTS = 0x02, MSG_CTS = 0x04, MSG_INLINE = 0x08, MSG_DONE = 0x10 };

#pragma pack(push, 1)
struct MsgHeader {
    uint8_t type;
    uint8_t credits;
    uint16_t pad;
    uint32_t tot_len;      // length of the text that follows
    uint64_t src_req;      // request cookie (0 on requests, echoed on acks)
    uint64_t payload_len;  // chunk bytes that follow the text (acks only)
};
#pragma pack(pop)

class Connection;

// ---- provider side ------------------------------------------------------------------------------------
class ShuffleServer {
public:
    using RequestHandler = std::function<void(const FetchRequest&, std::shared_ptr<Connection>)>;
    ShuffleServer(int port, RequestHandler on_request);
    ~ShuffleServer();
    void start();            // bind + listen(128) on INADDR_ANY:port, accept thread
    void stop();
    int port() const { return port_; }  // actual port (when constructed with 0)
    // ack + chunk; returns false when the connection is gone
    static bool reply(const std::shared_ptr<Connection>& c, uint64_t src_req, const std::string& ack_text,
                      const uint8_t* payload, size_t payload_len);

private:
    void accept_loop();
    void serve(std::shared_ptr<Connection> c);
    int port_;
    std::atomic<int> listen_fd_{-1};  // read by the accept thread, closed by stop()
    RequestHandler on_request_;
    std::atomic<bool> running_{false};
    std::thread accept_thread_;
    std::mutex mu_;
    std::vector<std::shared_ptr<Connection>> conns_;
    std::vector<std::thread> conn_threads_;
};

// ---- consumer side ------------------------------------------------------------------------------------
// Abstract fetcher (the reference's InputClient, src/Merger/InputClient.h:30-56): the merger only knows
// "fetch the next chunk of this map output into this buffer and tell me when it is there".
class InputClient {
public:
    using Done = std::function<void(bool ok, const FetchAck& ack)>;
    virtual ~InputClient() = default;
    virtual void start() = 0;
    virtual void stop() = 0;
    virtual void fetch(const std::string& host, FetchRequest req, uint8_t* dest, size_t capacity, Done done) = 0;
};

// Benchmark / test hook: when set, the consumer obtains its fetcher from this factory instead of creating a
// TcpInputClient (tests/native/cpu_merge_bench.cc feeds the consumer from memory the way baseline/ref_harness.cc feeds
// the reference's reducer, so the two CPU mergers can be compared without a network or disks in the way).
using InputClientFactory = std::function<std::unique_ptr<InputClient>(int port, int wqes_per_conn)>;
void set_input_client_factory(InputClientFactory f);
std::unique_ptr<InputClient> make_input_client(int port, int wqes_per_conn);

class TcpInputClient : public InputClient {
public:
    TcpInputClient(int port, int wqes_per_conn);
    ~TcpInputClient() override;
    void start() override {}
    void stop() override;
    void fetch(const std::string& host, FetchRequest req, uint8_t* dest, size_t capacity, Done done) override;

private:
    struct Pending { uint8_t* dest; size_t cap; Done done; };
    struct Peer {
        std::shared_ptr<Connection> conn;
        std::thread reader;
        std::mutex mu;
        std::map<uint64_t, Pending> inflight;
        std::vector<std::pair<std::string, uint64_t>> backlog;  // (request text, cookie) waiting for a credit
        int credits = 0;
        bool bad = false;
    };
    std::shared_ptr<Peer> peer_for(const std::string& host);
    void read_loop(std::shared_ptr<Peer> p);
    void fail_all(const std::shared_ptr<Peer>& p);
    int port_, wqes_;
    std::mutex mu_;
    std::map<std::string, std::shared_ptr<Peer>> peers_;
    std::atomic<uint64_t> next_cookie_{1};
    std::atomic<bool> stopping_{false};
};

}  // namespace uda
