// TEST INFRASTRUCTURE (oracle build shim) — not product code.
// torch ships fmt's headers but not fmt's os.cc, so the out-of-line pieces of fmt::file / fmt::ostream that the
// reference's write_model needs (fmt::output_file, src/plssvm/csvm.cpp:133) are provided here with plain POSIX calls.
#include "fmt/os.h"

#include <fcntl.h>
#include <unistd.h>

#include <cerrno>
#include <system_error>

namespace fmt {
inline namespace v12 {

file::file(cstring_view path, int oflag) {
    fd_ = ::open(path.c_str(), oflag, 0666);
    if (fd_ == -1) { throw std::system_error(errno, std::generic_category(), "cannot open file"); }
}
file::~file() noexcept {
    if (fd_ != -1) { ::close(fd_); }
}
void file::close() {
    if (fd_ == -1) { return; }
    const int r = ::close(fd_);
    fd_ = -1;
    if (r != 0) { throw std::system_error(errno, std::generic_category(), "cannot close file"); }
}
auto file::write(const void *buffer, size_t count) -> size_t {
    const char *p = static_cast<const char *>(buffer);
    size_t done = 0;
    while (done < count) {
        const ssize_t r = ::write(fd_, p + done, count - done);
        if (r < 0) {
            if (errno == EINTR) { continue; }
            throw std::system_error(errno, std::generic_category(), "cannot write to file");
        }
        done += static_cast<size_t>(r);
    }
    return done;
}

void ostream::grow(buffer<char> &buf, size_t) {
    if (buf.size() == buf.capacity()) { static_cast<ostream &>(buf).flush(); }
}
ostream::ostream(cstring_view path, const detail::ostream_params &params) :
    buffer<char>(grow), file_(path, params.oflag) {
    set(new char[params.buffer_size], params.buffer_size);
}
ostream::ostream(ostream &&other) noexcept :
    buffer<char>(grow, other.data(), other.size(), other.capacity()), file_(std::move(other.file_)) {
    other.clear();
    other.set(nullptr, 0);
}
ostream::~ostream() {
    flush();
    delete[] data();
}

}  // namespace v12
}  // namespace fmt
