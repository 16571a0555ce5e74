// column.cpp -- cylon::Column: device-resident Arrow-layout array (include/cylon/column.hpp).
#include "cylon/column.hpp"

#include <vector>

#include "util.hpp"

namespace cylon {

Column::~Column() {
  if (!owns_) return;
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (desc_.validity) cylb::pool_free(const_cast<uint8_t *>(desc_.validity), s);
  if (desc_.offsets) cylb::pool_free(const_cast<int32_t *>(desc_.offsets), s);
  if (desc_.data) cylb::pool_free(const_cast<void *>(desc_.data), s);
}

void Column::drop_validity() {
  if (owns_ && desc_.validity) cylb::pool_free(const_cast<uint8_t *>(desc_.validity), static_cast<cudaStream_t>(stream_));
  desc_.validity = nullptr;
  desc_.null_count = 0;
}

Status Column::Allocate(const std::string &name, int32_t type, int32_t byte_width, int64_t length, int64_t data_bytes, bool with_validity,
                        void *stream, std::shared_ptr<Column> *out) {
  auto c = std::make_shared<Column>();
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  c->name_ = name;
  c->owns_ = true;
  c->stream_ = stream;
  c->desc_.type = type;
  c->desc_.byte_width = (type == CYL_STRING || type == CYL_BINARY) ? 0 : byte_width;
  c->desc_.length = length;
  c->desc_.null_count = with_validity ? -1 : 0;  // -1 = unknown, as Arrow allows
  c->data_bytes_ = data_bytes;
  void *p = nullptr;
  CUDA_RETURN_IF_FAILED(cylb::pool_malloc(&p, data_bytes > 0 ? (size_t)data_bytes : 16, s));
  c->desc_.data = p;
  if (type == CYL_STRING || type == CYL_BINARY) {
    CUDA_RETURN_IF_FAILED(cylb::pool_malloc(&p, (size_t)(length + 1) * sizeof(int32_t), s));
    c->desc_.offsets = static_cast<int32_t *>(p);
  }
  if (with_validity) {
    CUDA_RETURN_IF_FAILED(cylb::pool_malloc(&p, detail::validity_bytes(length), s));
    c->desc_.validity = static_cast<uint8_t *>(p);
  }
  *out = c;
  return Status::OK();
}

Status Column::FromHost(const std::string &name, const cyl_column_t &h, void *stream, std::shared_ptr<Column> *out) {
  const bool var = h.type == CYL_STRING || h.type == CYL_BINARY;
  int64_t data_bytes = 0;
  if (var) {
    if (h.length > 0 && h.offsets == nullptr) return Status(Code::Invalid, "variable-width column without offsets");
    data_bytes = h.length > 0 ? (int64_t)h.offsets[h.length] - h.offsets[0] : 0;
  } else {
    if (h.byte_width <= 0) return Status(Code::Invalid, "fixed-width column needs byte_width > 0");
    data_bytes = h.length * (int64_t)h.byte_width;
  }
  std::shared_ptr<Column> c;
  RETURN_CYLON_STATUS_IF_FAILED(Allocate(name, h.type, h.byte_width, h.length, data_bytes, h.validity != nullptr, stream, &c));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (var) {
    const uint8_t *src = static_cast<const uint8_t *>(h.data) + (h.length > 0 ? h.offsets[0] : 0);
    if (data_bytes > 0) CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(c->mutable_data(), src, (size_t)data_bytes, cudaMemcpyHostToDevice, s));
    if (h.length > 0 && h.offsets[0] == 0) {
      CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(c->mutable_offsets(), h.offsets, (size_t)(h.length + 1) * 4, cudaMemcpyHostToDevice, s));
    } else {
      std::vector<int32_t> rebased((size_t)h.length + 1, 0);
      for (int64_t i = 0; i <= h.length && h.length > 0; i++) rebased[i] = h.offsets[i] - h.offsets[0];
      CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(c->mutable_offsets(), rebased.data(), rebased.size() * 4, cudaMemcpyHostToDevice, s));
      CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
    }
  } else if (data_bytes > 0) {
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(c->mutable_data(), h.data, (size_t)data_bytes, cudaMemcpyHostToDevice, s));
  }
  if (h.validity) {
    CUDA_RETURN_IF_FAILED(cudaMemsetAsync(c->mutable_validity(), 0, detail::validity_bytes(h.length), s));
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(c->mutable_validity(), h.validity, (size_t)((h.length + 7) / 8), cudaMemcpyHostToDevice, s));
    c->desc_.null_count = h.null_count;
  }
  *out = c;
  return Status::OK();
}

std::shared_ptr<Column> Column::View(const std::string &name, const cyl_column_t &d) {
  auto c = std::make_shared<Column>();
  c->name_ = name;
  c->desc_ = d;
  c->owns_ = false;
  if (d.type == CYL_STRING || d.type == CYL_BINARY) {
    int32_t ends[2] = {0, 0};
    if (d.length > 0 && d.offsets) {
      cudaMemcpy(&ends[0], d.offsets, 4, cudaMemcpyDeviceToHost);
      cudaMemcpy(&ends[1], d.offsets + d.length, 4, cudaMemcpyDeviceToHost);
    }
    c->data_bytes_ = ends[1] - ends[0];
  } else {
    c->data_bytes_ = d.length * (int64_t)d.byte_width;
  }
  return c;
}

Status Column::ToHost(void *data, int32_t *offsets, uint8_t *validity, void *stream) const {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (offsets && desc_.offsets) CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(offsets, desc_.offsets, (size_t)(desc_.length + 1) * 4, cudaMemcpyDeviceToHost, s));
  if (data && data_bytes_ > 0) CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(data, desc_.data, (size_t)data_bytes_, cudaMemcpyDeviceToHost, s));
  if (validity && desc_.validity)
    CUDA_RETURN_IF_FAILED(cudaMemcpyAsync(validity, desc_.validity, (size_t)((desc_.length + 7) / 8), cudaMemcpyDeviceToHost, s));
  CUDA_RETURN_IF_FAILED(cudaStreamSynchronize(s));
  return Status::OK();
}

}  // namespace cylon
