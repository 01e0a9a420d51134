/* Type-only stand-in for <infiniband/verbs.h>, written for the reference arm of bench.py.
 *
 * The reference's Merger headers reach RDMAComm.h, whose structs embed verbs types. The image has no
 * libibverbs, and the reference arm replaces the RDMA transport by a memory-fed InputClient
 * (baseline/ref_harness.cc), so only the *names and sizes* matter: nothing here is ever called. */
#ifndef UDA_REF_STUB_VERBS_H
#define UDA_REF_STUB_VERBS_H
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>   /* the real header pulls these in; the reference relies on that */
#ifdef __cplusplus
extern "C" {
#endif
enum ibv_access_flags { IBV_ACCESS_LOCAL_WRITE = 1, IBV_ACCESS_REMOTE_WRITE = 2, IBV_ACCESS_REMOTE_READ = 4 };
enum ibv_send_flags { IBV_SEND_SIGNALED = 2 };
enum ibv_wr_opcode { IBV_WR_RDMA_WRITE = 0, IBV_WR_SEND = 2 };
struct ibv_context { int async_fd; };
struct ibv_pd { struct ibv_context* context; };
struct ibv_comp_channel { struct ibv_context* context; int fd; };
struct ibv_cq { struct ibv_context* context; };
struct ibv_qp { struct ibv_context* context; uint32_t qp_num; };
struct ibv_mr { struct ibv_context* context; void* addr; uint64_t length; uint32_t lkey, rkey; };
struct ibv_sge { uint64_t addr; uint32_t length; uint32_t lkey; };
struct ibv_recv_wr { uint64_t wr_id; struct ibv_recv_wr* next; struct ibv_sge* sg_list; int num_sge; };
struct ibv_send_wr {
    uint64_t wr_id; struct ibv_send_wr* next; struct ibv_sge* sg_list; int num_sge; int opcode; int send_flags;
    struct { struct { uint64_t remote_addr; uint32_t rkey; } rdma; } wr;
};
struct ibv_wc { uint64_t wr_id; int status; int opcode; uint32_t byte_len; uint32_t qp_num; };
#ifdef __cplusplus
}
#endif
#endif
