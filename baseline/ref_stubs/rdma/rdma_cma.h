/* Type-only stand-in for <rdma/rdma_cma.h> (see infiniband/verbs.h beside it). */
#ifndef UDA_REF_STUB_RDMA_CMA_H
#define UDA_REF_STUB_RDMA_CMA_H
#include <infiniband/verbs.h>
#include <sys/socket.h>
#include <netinet/in.h>
#ifdef __cplusplus
extern "C" {
#endif
struct rdma_event_channel { int fd; };
struct rdma_cm_id { struct ibv_context* verbs; struct rdma_event_channel* channel; void* context; struct ibv_qp* qp; };
struct rdma_cm_event { struct rdma_cm_id* id; struct rdma_cm_id* listen_id; int event; int status; };
#ifdef __cplusplus
}
#endif
#endif
