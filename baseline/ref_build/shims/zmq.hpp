// zmq.hpp shim — the slice of cppzmq (4.2-era API) that the reference's io/ channel uses:
// context_t, socket_t (bind / connect / setsockopt / send / recv), message_t, error_t.
#ifndef MINIPS_SHIM_ZMQ_HPP_
#define MINIPS_SHIM_ZMQ_HPP_
#include <zmq.h>

#include <cerrno>
#include <cstring>
#include <exception>
#include <string>

namespace zmq {

class error_t : public std::exception {
 public:
  error_t() : errnum_(zmq_errno()) {}
  const char* what() const noexcept override { return zmq_strerror(errnum_); }
  int num() const { return errnum_; }

 private:
  int errnum_;
};

class message_t {
 public:
  message_t() { zmq_msg_init(&msg_); }
  explicit message_t(size_t size) {
    if (zmq_msg_init_size(&msg_, size) != 0) throw error_t();
  }
  message_t(const void* data, size_t size) {
    if (zmq_msg_init_size(&msg_, size) != 0) throw error_t();
    if (size) std::memcpy(zmq_msg_data(&msg_), data, size);
  }
  message_t(const message_t&) = delete;
  message_t& operator=(const message_t&) = delete;
  message_t(message_t&& o) noexcept {
    zmq_msg_init(&msg_);
    zmq_msg_move(&msg_, &o.msg_);
  }
  ~message_t() { zmq_msg_close(&msg_); }
  void rebuild() {
    zmq_msg_close(&msg_);
    zmq_msg_init(&msg_);
  }
  void rebuild(size_t size) {
    zmq_msg_close(&msg_);
    if (zmq_msg_init_size(&msg_, size) != 0) throw error_t();
  }
  void* data() { return zmq_msg_data(&msg_); }
  const void* data() const { return zmq_msg_data(const_cast<zmq_msg_t*>(&msg_)); }
  template <typename T>
  T* data() { return static_cast<T*>(data()); }
  size_t size() const { return zmq_msg_size(&msg_); }
  bool more() const { return zmq_msg_more(&msg_) != 0; }
  zmq_msg_t* handle() { return &msg_; }

 private:
  zmq_msg_t msg_;
};

class context_t {
 public:
  context_t() : ptr_(zmq_ctx_new()) {}
  explicit context_t(int io_threads, int max_sockets = 1023) : ptr_(zmq_ctx_new()) {
    zmq_ctx_set(ptr_, ZMQ_IO_THREADS, io_threads);
    zmq_ctx_set(ptr_, ZMQ_MAX_SOCKETS, max_sockets);
  }
  context_t(const context_t&) = delete;
  ~context_t() { close(); }
  void close() {
    if (ptr_) zmq_ctx_term(ptr_);
    ptr_ = nullptr;
  }
  operator void*() { return ptr_; }
  void* handle() { return ptr_; }

 private:
  void* ptr_;
};

class socket_t {
 public:
  socket_t(context_t& ctx, int type) : ptr_(zmq_socket(ctx.handle(), type)) {
    if (!ptr_) throw error_t();
  }
  socket_t(const socket_t&) = delete;
  ~socket_t() { close(); }
  void close() {
    if (ptr_) zmq_close(ptr_);
    ptr_ = nullptr;
  }
  operator void*() { return ptr_; }
  void setsockopt(int option, const void* optval, size_t len) {
    if (zmq_setsockopt(ptr_, option, optval, len) != 0) throw error_t();
  }
  template <typename T>
  void setsockopt(int option, const T& v) { setsockopt(option, &v, sizeof(T)); }
  void getsockopt(int option, void* optval, size_t* len) {
    if (zmq_getsockopt(ptr_, option, optval, len) != 0) throw error_t();
  }
  void bind(const std::string& addr) { bind(addr.c_str()); }
  void bind(const char* addr) {
    if (zmq_bind(ptr_, addr) != 0) throw error_t();
  }
  void connect(const std::string& addr) { connect(addr.c_str()); }
  void connect(const char* addr) {
    if (zmq_connect(ptr_, addr) != 0) throw error_t();
  }
  size_t send(const void* buf, size_t len, int flags = 0) {
    int n = zmq_send(ptr_, buf, len, flags);
    if (n >= 0) return static_cast<size_t>(n);
    if (zmq_errno() == EAGAIN) return 0;
    throw error_t();
  }
  bool send(message_t& msg, int flags = 0) {
    int n = zmq_msg_send(msg.handle(), ptr_, flags);
    if (n >= 0) return true;
    if (zmq_errno() == EAGAIN) return false;
    throw error_t();
  }
  size_t recv(void* buf, size_t len, int flags = 0) {
    int n = zmq_recv(ptr_, buf, len, flags);
    if (n >= 0) return static_cast<size_t>(n);
    if (zmq_errno() == EAGAIN) return 0;
    throw error_t();
  }
  bool recv(message_t* msg, int flags = 0) {
    int n = zmq_msg_recv(msg->handle(), ptr_, flags);
    if (n >= 0) return true;
    if (zmq_errno() == EAGAIN) return false;
    throw error_t();
  }

 private:
  void* ptr_;
};

}  // namespace zmq
#endif
