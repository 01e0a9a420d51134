#include "protocol.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "uda_rt.h"

namespace uda {

const char* const kPathTooLong = "MOF_PATH_SIZE_TOO_LONG";
const char* const kFetchRejected = "UDA_FETCH_REJECTED";

// ---------------------------------------------------------------------------------------------- argv
namespace {
bool to_long(const std::string& s, long* out) {
    errno = 0;
    char* end = nullptr;
    long v = strtol(s.c_str(), &end, 10);
    if (errno != 0) return false;
    *out = v;  // like strtol in the reference: trailing junk is ignored, "abc" parses as 0
    return true;
}
}  // namespace

bool parse_options(const std::vector<std::string>& argv, Options* op) {
    // getopt semantics: argv[0] is the program name and is skipped. The Java side passes no program
    // name ("-w", "256", "-r", ...), so the reference silently drops "-w" and treats its value as a stray
    // non-option; that behaviour is kept.
    static const struct { const char* lng; char shrt; } kLong[] = {
        {"wqesperconn", 'w'}, {"dataport", 'r'}, {"merge", 'a'}, {"mode", 'm'}, {"log", 'g'}, {"trace-level", 't'}, {"rdmabufsize", 's'}};
    for (size_t i = 1; i < argv.size(); ++i) {
        const std::string& a = argv[i];
        char opt = 0;
        std::string val;
        bool have_val = false;
        if (a.size() >= 3 && a[0] == '-' && a[1] == '-') {
            std::string name = a.substr(2);
            size_t eq = name.find('=');
            if (eq != std::string::npos) { val = name.substr(eq + 1); name = name.substr(0, eq); have_val = true; }
            for (auto& l : kLong) if (name == l.lng) opt = l.shrt;
            if (!opt) { printf("got invalid command line option %s\n", a.c_str()); continue; }
        } else if (a.size() >= 2 && a[0] == '-' && a != "--") {
            opt = a[1];
            if (a.size() > 2) { val = a.substr(2); have_val = true; }
        } else {
            continue;  // non-option argument
        }
        if (!strchr("wramgtbs", opt)) { printf("got invalid command line option choice=%c [%d] \n", opt, opt); continue; }
        if (!have_val) {
            if (i + 1 >= argv.size()) return false;
            val = argv[++i];
        }
        long v = 0;
        switch (opt) {
            case 'w': if (!to_long(val, &v)) return false; op->wqes_per_conn = (int)v; break;
            case 'r': if (!to_long(val, &v)) return false; op->data_port = (int)v; break;
            case 'a': if (!to_long(val, &v)) return false; op->merge_approach = (int)v; break;
            case 'm': if (!to_long(val, &v)) return false; op->mode = (int)v; break;
            case 'g': op->log_dir = val; break;
            case 't': if (!to_long(val, &v)) return false; op->log_threshold = (int)v; break;
            case 's': {
                if (!to_long(val, &v)) return false;
                long bytes = v * 1024;
                op->buf_size = bytes > 4096 ? (int)((bytes >> 12) << 12) : 4096;  // 4 KiB aligned, min 4 KiB
                break;
            }
            case 'b': break;  // accepted and ignored, as in the reference's getopt string
        }
    }
    return true;
}

// ---------------------------------------------------------------------------------------------- commands
bool parse_hadoop_cmd(const std::string& cmd, HadoopCmd* out) {
    out->params.clear();
    if (cmd.empty()) {  // empty string == EXIT
        out->count = 1;
        out->header = EXIT_MSG;
        return true;
    }
    size_t end = cmd.find(':');
    if (end == std::string::npos) return false;
    out->count = atoi(cmd.substr(0, end).c_str());
    size_t start = end + 1;
    end = cmd.find(':', start);
    if (end == std::string::npos) {
        out->header = atoi(cmd.substr(start).c_str());
        return true;
    }
    out->header = atoi(cmd.substr(start, end - start).c_str());
    start = end + 1;
    // count-2 params are ':'-delimited; the last one takes everything that is left (paths with ':')
    for (int i = 0; i < out->count - 2; ++i) {
        end = cmd.find(':', start);
        if (end == std::string::npos) return false;
        out->params.push_back(cmd.substr(start, end - start));
        start = end + 1;
    }
    out->params.push_back(cmd.substr(start));
    return true;
}

std::string form_cmd(int cmd, const std::vector<std::string>& params) {
    std::string s = std::to_string(params.size() + 1) + ":" + std::to_string(cmd);
    for (auto& p : params) { s += ':'; s += p; }
    return s;
}

bool parse_init_params(const HadoopCmd& c, InitParams* p) {
    if (c.header != INIT_MSG || c.params.size() < 10) return false;
    const auto& a = c.params;
    p->num_maps = atoi(a[0].c_str());
    p->job_id = a[1];
    p->reduce_task_id = a[2];
    p->lpq_size = atoi(a[3].c_str());
    p->max_buffer_bytes = atoll(a[4].c_str());
    p->min_buffer_bytes = atoll(a[5].c_str());
    p->key_class = a[6];
    p->codec_class = a[7];
    p->codec_block_size = atoi(a[8].c_str());
    p->shuffle_memory = atoll(a[9].c_str());
    p->local_dirs.clear();
    if (a.size() > 10) {
        int nd = atoi(a[10].c_str());
        if (nd > 0 && (size_t)(11 + nd) <= a.size())
            for (int i = 0; i < nd; ++i) p->local_dirs.push_back(a[11 + i]);
    }
    return true;
}

// ---------------------------------------------------------------------------------------------- fetch
std::string format_fetch_request(const FetchRequest& r) {
    char buf[kMaxMsgText + 64];
    int n = snprintf(buf, sizeof(buf), "%s:%s:%lld:%d:%lu:%lu:%d:%lld:%s:%lld:%lld", r.job_id.c_str(), r.map_id.c_str(),
                     r.map_offset, r.reduce_id, (unsigned long)r.remote_addr, (unsigned long)r.request_ptr, r.chunk_size,
                     r.file_offset, r.path.c_str(), r.raw_length, r.part_length);
    if (n < 0 || (size_t)n >= kMaxMsgText) throw UdaError("trying to fetch a message too big");
    return std::string(buf, (size_t)n);
}

namespace {
bool next_field(const std::string& s, size_t* pos, std::string* out) {
    size_t end = s.find(':', *pos);
    if (end == std::string::npos) return false;
    *out = s.substr(*pos, end - *pos);
    *pos = end + 1;
    return true;
}
}  // namespace

bool parse_fetch_request(const std::string& s, FetchRequest* r) {
    size_t pos = 0;
    std::string f;
    if (!next_field(s, &pos, &r->job_id)) return false;
    if (!next_field(s, &pos, &r->map_id)) return false;
    if (!next_field(s, &pos, &f)) return false;
    r->map_offset = atoll(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    r->reduce_id = atoi(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    r->remote_addr = strtoull(f.c_str(), nullptr, 10);
    if (!next_field(s, &pos, &f)) return false;
    r->request_ptr = strtoull(f.c_str(), nullptr, 10);
    if (!next_field(s, &pos, &f)) return false;
    r->chunk_size = atoi(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    r->file_offset = atoll(f.c_str());
    if (!next_field(s, &pos, &r->path)) return false;
    if (!next_field(s, &pos, &f)) return false;
    r->raw_length = atoll(f.c_str());
    r->part_length = atoll(s.substr(pos).c_str());
    return true;
}

std::string format_fetch_ack(const FetchAck& a) {
    char buf[kMaxMsgText + 64];
    int n = snprintf(buf, sizeof(buf), "%lld:%lld:%lld:%lld:%s:", a.raw_length, a.part_length, a.sent, a.file_offset, a.path.c_str());
    if (n < 0 || (size_t)n >= kMaxMsgText) throw UdaError("trying to send a message too big");
    return std::string(buf, (size_t)n);
}

bool parse_fetch_ack(const std::string& s, FetchAck* a) {
    size_t pos = 0;
    std::string f;
    if (!next_field(s, &pos, &f)) return false;
    a->raw_length = atoll(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    a->part_length = atoll(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    a->sent = atoll(f.c_str());
    if (!next_field(s, &pos, &f)) return false;
    a->file_offset = atoll(f.c_str());
    if (!next_field(s, &pos, &a->path)) return false;
    return true;
}

}  // namespace uda
