// boost/iostreams/stream.hpp shim — boost::iostreams::stream<Device> for the reference's devices
// (general_source / general_sink / hdfs_seekable_device: read / write / close / good) built on a
// std::streambuf.  The stream's arguments construct the Device, like boost does.
#ifndef MINIPS_SHIM_BOOST_IOSTREAMS_STREAM_HPP_
#define MINIPS_SHIM_BOOST_IOSTREAMS_STREAM_HPP_
#include <ios>
#include <iostream>
#include <istream>
#include <ostream>
#include <streambuf>
#include <type_traits>
#include <utility>
#include <vector>
namespace boost {
namespace iostreams {
typedef std::streamoff stream_offset;
struct device_tag {};
struct multichar_tag {};
struct input {};
struct output {};
struct seekable {};
struct input_seekable : input {};
struct output_seekable : output {};
struct seekable_device_tag : device_tag, seekable {};
struct closable_tag {};
struct optimally_buffered_tag {};
struct source_tag : device_tag, input {};
struct sink_tag : device_tag, output {};

template <typename Device>
class device_buf : public std::streambuf {
 public:
  device_buf() : buf_(1 << 16), obuf_(1 << 16) {}
  Device& device() { return dev_; }
  void close_device() {
    sync();
    dev_.close();
  }
  ~device_buf() override { sync(); }

 protected:
  int_type underflow() override {
    if (gptr() < egptr()) return traits_type::to_int_type(*gptr());
    const std::streamsize n = dev_.read(buf_.data(), static_cast<std::streamsize>(buf_.size()));
    if (n <= 0) return traits_type::eof();
    setg(buf_.data(), buf_.data(), buf_.data() + n);
    return traits_type::to_int_type(*gptr());
  }
  int_type overflow(int_type c) override {
    if (pbase() == nullptr) setp(obuf_.data(), obuf_.data() + obuf_.size());
    flush_out();
    if (c != traits_type::eof()) {
      *pptr() = traits_type::to_char_type(c);
      pbump(1);
    }
    return traits_type::not_eof(c);
  }
  int sync() override {
    flush_out();
    return 0;
  }

 private:
  void flush_out() {
    if (pbase() && pptr() > pbase()) {
      dev_.write(pbase(), pptr() - pbase());
      setp(obuf_.data(), obuf_.data() + obuf_.size());
    }
  }
  Device dev_;
  std::vector<char> buf_, obuf_;
};

template <typename Device>
struct stream_base {
  typedef typename Device::category cat;
  typedef typename std::conditional<
      std::is_base_of<seekable, cat>::value, std::iostream,
      typename std::conditional<std::is_base_of<input, cat>::value, std::istream, std::ostream>::type>::type type;
};

template <typename Device>
class stream : public stream_base<Device>::type {
  typedef typename stream_base<Device>::type base_type;

 public:
  stream() : base_type(&buf_) {}
  template <typename... Args>
  explicit stream(Args&&... args) : base_type(&buf_) {
    open(std::forward<Args>(args)...);
  }
  ~stream() { buf_.pubsync(); }
  template <typename... Args>
  void open(Args&&... args) {
    buf_.device() = Device(std::forward<Args>(args)...);
    opened_ = true;
    if (!buf_.device().good()) this->setstate(std::ios_base::failbit);
  }
  bool is_open() const { return opened_; }
  void close() {
    if (opened_) buf_.close_device();
    opened_ = false;
  }
  Device& operator*() { return buf_.device(); }
  Device* operator->() { return &buf_.device(); }

 private:
  device_buf<Device> buf_;
  bool opened_ = false;
};
}  // namespace iostreams
}  // namespace boost
#endif
