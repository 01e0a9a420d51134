/* uda_b200 — C API of the host runtime (libuda.so): the provider ("MOFSupplier") and consumer
 * ("NetMerger") roles behind a handle, plus protocol helpers. The JNI exports for
 * com.mellanox.hadoop.mapred.UdaBridge live in the same library (csrc/jni/uda_bridge.cc) and call these.
 *
 * Reference surface: src/UdaBridge.cc:187-330 (startNative / doCommandNative / reduceExitMsgNative /
 * setLogLevelNative) and the six Java up-calls (UdaBridge.java:85-145). */
#ifndef UDA_RT_API_H
#define UDA_RT_API_H

#ifdef __cplusplus
extern "C" {
#endif

typedef void* uda_role_t;

/* Up-calls into the embedding application; any pointer may be NULL. */
typedef struct {
    void* user;
    void (*fetch_over_message)(void* user);
    void (*data_from_uda)(void* user, const void* buf, int len);
    /* return 1 and fill the outputs when the map output exists, else 0 */
    int (*get_path_uda)(void* user, const char* job, const char* map, int reduce, long long* start_offset,
                        long long* raw_length, long long* part_length, char* path_out, int path_cap);
    /* return 1 and write the value, or 0 to use the default */
    int (*get_conf_data)(void* user, const char* name, const char* dflt, char* out, int cap);
    void (*log_to_java)(void* user, const char* msg, int severity);
    void (*failure_in_uda)(void* user);
} uda_host_callbacks_t;

/* argv as Java passes it: "-w 256 -r 9011 -a 1 -m 1 -g <logdir> -s 1024" (no program name). */
uda_role_t uda_rt_start(int is_net_merger, int argc, const char* const* argv, int log_level, int log_to_unique_file,
                        const uda_host_callbacks_t* cb);
int uda_rt_do_command(uda_role_t h, const char* cmd); /* 0 ok / ignored-after-failure, -1 failure (see last_error) */
void uda_rt_reduce_exit(uda_role_t h);
void uda_rt_set_log_level(int level);
int uda_rt_failed(uda_role_t h);
const char* uda_rt_last_error(uda_role_t h);
int uda_rt_port(uda_role_t h);
long long uda_rt_stat(uda_role_t h, const char* what);
void uda_rt_destroy(uda_role_t h);
const char* uda_rt_version(void);

/* protocol helpers (golden tests) */
int uda_rt_vint_encode(long long v, unsigned char* out9);
int uda_rt_vint_decode(const unsigned char* p, int n, long long* out);
int uda_rt_vint_size(long long v);
int uda_rt_parse_cmd(const char* s, char* out, int cap);
int uda_rt_form_cmd(int cmd, int nparams, const char* const* params, char* out, int cap);
int uda_rt_parse_options(int argc, const char* const* argv, long long* out7, char* log_dir, int cap);
int uda_rt_fetch_request_roundtrip(const char* in, char* out, int cap);
int uda_rt_fetch_ack_roundtrip(const char* in, char* out, int cap);
int uda_rt_compare_keys(int kind, const unsigned char* a, int la, const unsigned char* b, int lb);
int uda_rt_key_kind(const char* java_class);
int uda_rt_snappy_uncompress(const unsigned char* in, long long n, unsigned char* out, long long out_len);
long long uda_rt_snappy_store(const unsigned char* in, long long n, unsigned char* out, long long cap);
/* greedy Snappy compressor with real copy elements (tests / benchmarks; cap >= 32 + n + n/6 is always enough) */
long long uda_rt_snappy_compress(const unsigned char* in, long long n, unsigned char* out, long long cap);
/* built-in LZO1X block codec: greedy compressor (cap >= n + n/16 + 64 + 3 is always enough) and decoder (0 = ok) */
long long uda_rt_lzo1x_compress(const unsigned char* in, long long n, unsigned char* out, long long cap);
int uda_rt_lzo1x_decompress(const unsigned char* in, long long n, unsigned char* out, long long out_len);
/* same decoder with the device kernel's warp-cooperative copy emulated lane by lane on the host (tests) */
int uda_rt_lzo1x_decompress_lanes(const unsigned char* in, long long n, unsigned char* out, long long out_len, int lanes);
int uda_rt_plan_memory(int num_maps, long long max_buf, long long min_buf, long long shuffle_mem, int compressed,
                       double ratio, long long* out4);
int uda_rt_selftest_queues(void);
/* test hook: the merge's record parser over a segment arriving in blocks of `chunk` bytes; records before the EOF marker or -1 */
long long uda_rt_selftest_parse_segment(const unsigned char* data, long long n, int chunk, unsigned long long* digest);
/* thread wrapper (exception -> one failure up-call) and Fisher-Yates fetch-order helpers; 0 = all invariants held */
int uda_rt_selftest_threads(unsigned long long seed, int n, int* perm_out);

#ifdef __cplusplus
}
#endif
#endif
