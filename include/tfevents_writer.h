// tfevents_writer.h — TensorBoard event-file writer for the two CLIs.  Mirrors the interface the
// reference mains use from the vendored logger (libs/tensorboard_logger/include/tensorboard_logger.h:50-161:
// `TensorBoardLogger(path)`, `add_scalar(tag, step, value)`, `add_text(tag, step, text)`; call sites
// src/ppo_continuous_action.cpp:281-282,426-432,575-583,621-624) and produces the same on-disk format, without protobuf:
//   record  = u64 length | u32 masked_crc32c(length) | Event bytes | u32 masked_crc32c(Event bytes)
//             (libs/tensorboard_logger/src/tensorboard_logger.cc:314-327, masking rule src/crc.cc:235-238)
//   Event   = { 1: double wall_time, 2: int64 step, 3: string file_version | 5: Summary }
//   Summary = { 1: repeated Value }, Value = { 1: tag, 2: float simple_value | 8: TensorProto, 9: SummaryMetadata }
//   text    = TensorProto{ 1: dtype DT_STRING(7), 8: string_val } + SummaryMetadata{ 1: PluginData{ 1: "text" } }
//             (tensorboard_logger.cc:164-184)
// The message bytes are hand-encoded protobuf wire format (varint / fixed32 / fixed64 / length-delimited).  The file
// name must contain "tfevents" (tensorboard_logger.h:56-60); the file is truncated unless `resume` is set (:63-66).
// Unlike the reference there is no flusher thread: flush() is called by the CLI once per iteration.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <stdexcept>
#include <string>

namespace tfevents {

inline uint32_t crc32c(const void *data, size_t n) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; i++) {
      uint32_t c = i;
      for (int k = 0; k < 8; k++) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : c >> 1;   // Castagnoli, reflected
      table[i] = c;
    }
    init = true;
  }
  uint32_t crc = 0xFFFFFFFFu;
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
  return ~crc;
}
inline uint32_t masked_crc32c(const void *data, size_t n) {
  const uint32_t crc = crc32c(data, n);
  return ((crc >> 15) | (crc << 17)) + 0xa282ead8u;
}

// protobuf wire-format helpers
inline void put_varint(std::string &b, uint64_t v) {
  while (v >= 0x80u) { b.push_back((char)((v & 0x7Fu) | 0x80u)); v >>= 7; }
  b.push_back((char)v);
}
inline void put_key(std::string &b, int field, int wire) { put_varint(b, ((uint64_t)field << 3) | (uint64_t)wire); }
inline void put_bytes(std::string &b, int field, const std::string &s) { put_key(b, field, 2); put_varint(b, s.size()); b += s; }
inline void put_double(std::string &b, int field, double v) { put_key(b, field, 1); char t[8]; std::memcpy(t, &v, 8); b.append(t, 8); }
inline void put_float(std::string &b, int field, float v) { put_key(b, field, 5); char t[4]; std::memcpy(t, &v, 4); b.append(t, 4); }
inline void put_int(std::string &b, int field, int64_t v) { put_key(b, field, 0); put_varint(b, (uint64_t)v); }

class Writer {
 public:
  explicit Writer(const std::string &log_file, bool resume = false) {
    const size_t slash = log_file.find_last_of('/');
    const std::string base = slash == std::string::npos ? log_file : log_file.substr(slash + 1);
    if (base.find("tfevents") == std::string::npos)
      throw std::runtime_error("A valid event file must contain substring \"tfevents\" in its basename, got " + base);
    f_ = std::fopen(log_file.c_str(), resume ? "ab" : "wb");
    if (!f_) throw std::runtime_error("failed to open log_file " + log_file);
    if (!resume) {  // TensorFlow's writers open every file with this record; readers use it to detect the format version
      std::string ev;
      put_double(ev, 1, (double)std::time(nullptr));
      put_bytes(ev, 3, "brain.Event:2");
      write_record(ev);
    }
  }
  ~Writer() { if (f_) std::fclose(f_); }
  Writer(const Writer &) = delete;
  Writer &operator=(const Writer &) = delete;

  int add_scalar(const std::string &tag, int64_t step, double value) { return add_scalar(tag, step, (float)value); }
  int add_scalar(const std::string &tag, int64_t step, float value) {
    std::string val;
    put_bytes(val, 1, tag);
    put_float(val, 2, value);
    return add_event(step, val);
  }
  int add_text(const std::string &tag, int64_t step, const std::string &text) {
    std::string plugin, meta, tensor, val;
    put_bytes(plugin, 1, "text");
    put_bytes(meta, 1, plugin);
    put_int(tensor, 1, 7);          // DT_STRING
    put_bytes(tensor, 8, text);
    put_bytes(val, 1, tag);
    put_bytes(val, 8, tensor);
    put_bytes(val, 9, meta);
    return add_event(step, val);
  }
  void flush() { if (f_) std::fflush(f_); }

 private:
  int add_event(int64_t step, const std::string &value) {
    std::string summary, ev;
    put_bytes(summary, 1, value);
    put_double(ev, 1, (double)std::time(nullptr));   // tensorboard_logger.cc:305-311
    put_int(ev, 2, step);
    put_bytes(ev, 5, summary);
    return write_record(ev);
  }
  int write_record(const std::string &ev) {
    const uint64_t len = ev.size();
    const uint32_t len_crc = masked_crc32c(&len, sizeof(len)), data_crc = masked_crc32c(ev.data(), ev.size());
    bool ok = std::fwrite(&len, sizeof(len), 1, f_) == 1 && std::fwrite(&len_crc, sizeof(len_crc), 1, f_) == 1 &&
              std::fwrite(ev.data(), 1, ev.size(), f_) == ev.size() && std::fwrite(&data_crc, sizeof(data_crc), 1, f_) == 1;
    return ok ? 0 : -1;
  }
  std::FILE *f_ = nullptr;
};

}  // namespace tfevents
