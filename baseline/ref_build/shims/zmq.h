/* zmq.h shim — prototypes and constants of the libzmq 4.x C API that the reference uses.  The
 * implementation is pyzmq's bundled libzmq-*.so.5 (headers are not shipped with the wheel). */
#ifndef MINIPS_SHIM_ZMQ_H_
#define MINIPS_SHIM_ZMQ_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define ZMQ_VERSION_MAJOR 4
#define ZMQ_VERSION_MINOR 3
#define ZMQ_VERSION_PATCH 5
#define ZMQ_IO_THREADS 1
#define ZMQ_MAX_SOCKETS 2
#define ZMQ_PAIR 0
#define ZMQ_PUB 1
#define ZMQ_SUB 2
#define ZMQ_REQ 3
#define ZMQ_REP 4
#define ZMQ_DEALER 5
#define ZMQ_ROUTER 6
#define ZMQ_PULL 7
#define ZMQ_PUSH 8
#define ZMQ_IDENTITY 5
#define ZMQ_ROUTING_ID 5
#define ZMQ_RCVMORE 13
#define ZMQ_LINGER 17
#define ZMQ_SNDHWM 23
#define ZMQ_RCVHWM 24
#define ZMQ_RCVTIMEO 27
#define ZMQ_SNDTIMEO 28
#define ZMQ_MORE 1
#define ZMQ_DONTWAIT 1
#define ZMQ_NOBLOCK ZMQ_DONTWAIT
#define ZMQ_SNDMORE 2
#define ETERM 156384765
typedef struct zmq_msg_t {
  unsigned char _[64] __attribute__((aligned(sizeof(void*))));
} zmq_msg_t;
typedef void(zmq_free_fn)(void* data, void* hint);
int zmq_errno(void);
const char* zmq_strerror(int errnum);
void* zmq_ctx_new(void);
int zmq_ctx_term(void* context);
int zmq_ctx_destroy(void* context);
int zmq_ctx_shutdown(void* context);
int zmq_ctx_set(void* context, int option, int optval);
int zmq_ctx_get(void* context, int option);
void* zmq_init(int io_threads);
int zmq_term(void* context);
int zmq_msg_init(zmq_msg_t* msg);
int zmq_msg_init_size(zmq_msg_t* msg, size_t size);
int zmq_msg_init_data(zmq_msg_t* msg, void* data, size_t size, zmq_free_fn* ffn, void* hint);
int zmq_msg_send(zmq_msg_t* msg, void* s, int flags);
int zmq_msg_recv(zmq_msg_t* msg, void* s, int flags);
int zmq_msg_close(zmq_msg_t* msg);
int zmq_msg_move(zmq_msg_t* dest, zmq_msg_t* src);
int zmq_msg_copy(zmq_msg_t* dest, zmq_msg_t* src);
void* zmq_msg_data(zmq_msg_t* msg);
size_t zmq_msg_size(const zmq_msg_t* msg);
int zmq_msg_more(const zmq_msg_t* msg);
void* zmq_socket(void*, int type);
int zmq_close(void* s);
int zmq_setsockopt(void* s, int option, const void* optval, size_t optvallen);
int zmq_getsockopt(void* s, int option, void* optval, size_t* optvallen);
int zmq_bind(void* s, const char* addr);
int zmq_connect(void* s, const char* addr);
int zmq_send(void* s, const void* buf, size_t len, int flags);
int zmq_recv(void* s, void* buf, size_t len, int flags);
#ifdef __cplusplus
}
#endif
#endif
