// Text protocols spoken between the Java plug-ins and the native side, and between consumer and
// provider. Byte-compatible with the reference, including its quirks (SURVEY §2.7 b-d):
//   * argv of startNative            src/CommUtils/C2JNexus.cc:43-137
//   * "<n+1>:<cmd>:<p1>:...:<pn>"    src/CommUtils/C2JNexus.cc:152-207 (last param swallows the rest,
//                                    empty string == EXIT), formed by UdaCmd.formCmd (UdaPlugin.java:577-586)
//   * INIT parameters                src/Merger/reducer.cc:56-99
//   * fetch request (11 fields)      src/DataNet/RDMAClient.cc:559-600, parsed src/MOFServer/MOFServlet.cc:28-96
//   * fetch ack                      src/DataNet/RDMAServer.cc:537-631, parsed src/Merger/MergeManager.cc:367-409
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace uda {

// ---- argv ---------------------------------------------------------------------------------------------
enum Mode { STANDALONE = 0, INTEGRATED = 1 };
struct Options {
    int wqes_per_conn = 256;  // -w  (kept for compatibility: credits per connection)
    int data_port = 9011;     // -r
    int merge_approach = 1;   // -a  0 unimplemented, 1 online, 2 hybrid, 3 device merge (no counterpart in the reference)
    int mode = INTEGRATED;    // -m
    std::string log_dir = "default";  // -g
    int log_threshold = -1;   // -t  (-1: not given)
    int buf_size = 0;         // -s  KiB on the command line; stored in bytes, 4 KiB aligned, min 4 KiB
};
// Returns false on a malformed numeric argument (the reference prints the argv and fails start-up).
bool parse_options(const std::vector<std::string>& argv, Options* out);

// ---- commands -----------------------------------------------------------------------------------------
enum Cmd {
    EXIT_MSG = 0, NEW_MAP_MSG = 1, FINAL_MSG = 2, RESULT = 3, FETCH_MSG = 4, FETCH_OVER_MSG = 5, JOB_OVER_MSG = 6,
    INIT_MSG = 7, MORE_MSG = 8, NETLEV_REDUCE_LAUNCHED = 9
};
struct HadoopCmd {
    int count = 0;  // header + params
    int header = EXIT_MSG;
    std::vector<std::string> params;
};
bool parse_hadoop_cmd(const std::string& s, HadoopCmd* out);
std::string form_cmd(int cmd, const std::vector<std::string>& params);

// ---- INIT ---------------------------------------------------------------------------------------------
struct InitParams {
    int num_maps = 0;
    std::string job_id, reduce_task_id;
    int lpq_size = 0;
    long long max_buffer_bytes = 0, min_buffer_bytes = 0;
    std::string key_class, codec_class;  // codec "null" = none
    int codec_block_size = 0;
    long long shuffle_memory = 0;
    std::vector<std::string> local_dirs;
};
bool parse_init_params(const HadoopCmd& cmd, InitParams* out);

// ---- fetch request / ack ------------------------------------------------------------------------------
constexpr size_t kMaxMsgText = 800;   // NETLEV_FETCH_REQSIZE (src/include/NetlevComm.h:30)
constexpr size_t kMaxMofPath = 600;   // NETLEV_MOF_PATH_MAX_SIZE (src/include/NetlevComm.h:31)
extern const char* const kPathTooLong;  // "MOF_PATH_SIZE_TOO_LONG"
// path field of an ack that rejects a malformed or unservable request (no reference counterpart: the reference never
// answers such a request and the reducer waits for ever)
extern const char* const kFetchRejected;  // "UDA_FETCH_REJECTED"

struct FetchRequest {
    std::string job_id, map_id;
    long long map_offset = 0;  // bytes of this partition already fetched
    int reduce_id = 0;
    uint64_t remote_addr = 0;  // destination buffer token (an address in the reference)
    uint64_t request_ptr = 0;  // opaque cookie echoed in the ack
    int chunk_size = 0;
    long long file_offset = 0; // startOffset in file.out (0 / unknown on the first fetch)
    std::string path;          // empty on the first fetch
    long long raw_length = 0, part_length = 0;
};
std::string format_fetch_request(const FetchRequest& r);          // throws UdaError when >= 800 bytes
bool parse_fetch_request(const std::string& s, FetchRequest* out); // false when a ':' is missing

struct FetchAck {
    long long raw_length = 0, part_length = 0;
    long long sent = 0;        // bytes delivered by this chunk
    long long file_offset = 0;
    std::string path;
};
std::string format_fetch_ack(const FetchAck& a);
bool parse_fetch_ack(const std::string& s, FetchAck* out);

}  // namespace uda
