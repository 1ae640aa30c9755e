// cli_common.h — small host-side helpers shared by the two command-line drivers: `--name value` flag parsing in the
// style of the reference's args.hxx ValueFlags (long names only, bools as 0/1: src/ppo_continuous_action.cpp:199-224),
// the TicToc wall-clock timer (include/tictoc.h:11-38), a scalar logger that keeps the reference's TensorBoard tag
// names (src/ppo_continuous_action.cpp:426-432,575-583) and writes both the reference's TensorBoard event file
// (include/tfevents_writer.h) and JSON lines, and the rank/world bootstrap that replaces
// MPI_Init + MPI_Bcast(ncclUniqueId) (src/ac_ppo_continuous_action.cpp:301-303, src/distributed.cpp:66-79).
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <filesystem>
#include <string>
#include <vector>
#include <system_error>
#include <thread>
#include <optional>
#include <condition_variable>
#include <mutex>
#include <vector>

#include <memory>

#include "ppo_b200.h"
#include "tfevents_writer.h"

namespace cli {

class Flags {
 public:
  void add(const std::string &name, int *v, const std::string &help) { add_impl(name, help, [v](const char *s) { *v = std::atoi(s); }, std::to_string(*v)); }
  void add(const std::string &name, long *v, const std::string &help) { add_impl(name, help, [v](const char *s) { *v = std::atol(s); }, std::to_string(*v)); }
  void add(const std::string &name, float *v, const std::string &help) { add_impl(name, help, [v](const char *s) { *v = (float)std::atof(s); }, std::to_string(*v)); }
  void add(const std::string &name, bool *v, const std::string &help) { add_impl(name, help, [v](const char *s) { *v = std::atoi(s) != 0; }, *v ? "1" : "0"); }
  void add(const std::string &name, std::string *v, const std::string &help) { add_impl(name, help, [v](const char *s) { *v = s; }, *v); }
  // returns 0 to continue, 1 after --help (caller exits 0), 2 on a parse error (caller exits 1 like the reference)
  int parse(int argc, char **argv) {
    for (int i = 1; i < argc; i++) {
      std::string a = argv[i];
      if (a == "-h" || a == "--help") { print_help(argv[0]); return 1; }
      std::string name, value;
      bool have_value = false;
      if (a.rfind("--", 0) == 0) {
        size_t eq = a.find('=');
        name = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
        if (eq != std::string::npos) { value = a.substr(eq + 1); have_value = true; }
      }
      auto it = setters_.find(name);
      if (it == setters_.end()) { std::cerr << "Flag could not be matched: " << a << std::endl; print_help(argv[0]); return 2; }
      if (!have_value) {
        if (i + 1 >= argc) { std::cerr << "Flag '" << name << "' requires an argument" << std::endl; return 2; }
        value = argv[++i];
      }
      it->second(value.c_str());
    }
    return 0;
  }
  void print_help(const char *prog) const {
    std::cout << "  " << prog << " {OPTIONS}\n\n  OPTIONS:\n";
    for (const auto &h : help_) std::cout << "      --" << h << "\n";
  }

 private:
  void add_impl(const std::string &name, const std::string &help, std::function<void(const char *)> f, const std::string &def) {
    setters_[name] = std::move(f);
    help_.push_back(name + "  " + help + " (default " + def + ")");
  }
  std::map<std::string, std::function<void(const char *)>> setters_;
  std::vector<std::string> help_;
};

class TicToc {  // include/tictoc.h:11-38
 public:
  void tic() { t0_ = std::chrono::steady_clock::now(); }
  double tocvalue() const { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0_).count(); }
  void toc(const std::string &label) const { std::cout << label << " " << tocvalue() << " s" << std::endl; }

 private:
  std::chrono::steady_clock::time_point t0_ = std::chrono::steady_clock::now();
};

// Every scalar goes to the TensorBoard event file the reference writes (`<exp_folder>/tfevents_logs.pb`,
// src/ppo_continuous_action.cpp:281; same tags and steps) and, as one JSON object per scalar
// {"tag": ..., "step": ..., "value": ...}, to scalars.jsonl; text entries carry "text".
class ScalarLogger {
 public:
  explicit ScalarLogger(const std::string &path, const std::string &tfevents_path = "") : out_(path, std::ios::app) {
    if (!tfevents_path.empty()) tb_ = std::make_unique<tfevents::Writer>(tfevents_path);
  }
  void add_scalar(const std::string &tag, long step, double value) {
    if (tb_) tb_->add_scalar(tag, step, value);
    if (out_) out_ << "{\"tag\": \"" << tag << "\", \"step\": " << step << ", \"value\": " << value << "}\n";
  }
  void add_text(const std::string &tag, long step, const std::string &text) {
    if (tb_) tb_->add_text(tag, step, text);
    if (!out_) return;
    out_ << "{\"tag\": \"" << tag << "\", \"step\": " << step << ", \"text\": \"";
    for (char c : text) { if (c == '\n') out_ << "\\n"; else if (c == '"') out_ << "\\\""; else out_ << c; }
    out_ << "\"}\n";
  }
  void flush() { out_.flush(); if (tb_) tb_->flush(); }

 private:
  std::ofstream out_;
  std::unique_ptr<tfevents::Writer> tb_;
};

// Data-parallel bootstrap without MPI: PB_RANK / PB_WORLD (or torchrun's RANK / WORLD_SIZE / LOCAL_RANK) name the
// replica; rank 0 writes the 128-byte ncclUniqueId to PB_NCCL_ID_FILE, the other ranks wait for it.
struct Dist {
  int rank = 0, world = 1, local_rank = 0;
  static int env_int(const char *a, const char *b, int def) {
    const char *v = std::getenv(a);
    if (!v && b) v = std::getenv(b);
    return v ? std::atoi(v) : def;
  }
  static Dist from_env() {
    Dist d;
    d.rank = env_int("PB_RANK", "RANK", 0);
    d.world = env_int("PB_WORLD", "WORLD_SIZE", 1);
    d.local_rank = env_int("PB_LOCAL_RANK", "LOCAL_RANK", d.rank);
    return d;
  }
  // returns 0 on success; fills id128
  int exchange_unique_id(unsigned char *id128) const {
    const char *path = std::getenv("PB_NCCL_ID_FILE");
    if (!path) { std::cerr << "PB_NCCL_ID_FILE must be set when PB_WORLD > 1" << std::endl; return -1; }
    // Lifecycle: rank 0 removes whatever an earlier run left at the path before it publishes the new id, and the
    // other ranks ignore a file that is older than ten minutes (a leftover of a crashed run whose rank 0 never got here);
    // remove_unique_id() deletes the file once ncclCommInitRank - a collective - has returned, i.e. every rank has read it.
    if (rank == 0) {
      std::remove(path);
      if (pb_comm_unique_id(id128) != PB_OK) return -1;
      std::string tmp = std::string(path) + ".tmp";
      FILE *f = std::fopen(tmp.c_str(), "wb");
      if (!f) return -1;
      bool ok = std::fwrite(id128, 1, 128, f) == 128;
      std::fclose(f);
      return ok && std::rename(tmp.c_str(), path) == 0 ? 0 : -1;
    }
    for (int tries = 0; tries < 6000; tries++) {  // up to 60 s
      std::error_code ec;
      const auto mtime = std::filesystem::last_write_time(path, ec);
      const bool fresh = !ec && std::filesystem::file_time_type::clock::now() - mtime < std::chrono::minutes(10);
      FILE *f = fresh ? std::fopen(path, "rb") : nullptr;
      if (f) {
        size_t n = std::fread(id128, 1, 128, f);
        std::fclose(f);
        if (n == 128) return 0;
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    return -1;
  }
  void remove_unique_id() const {
    const char *path = std::getenv("PB_NCCL_ID_FILE");
    if (rank == 0 && path) std::remove(path);
  }
};

// model_latest_%09d.pth / optimizer_latest_%09d.pth / model_final.pth (src/ppo_continuous_action.cpp:173-180,546-563,587): the CLI dumps
// parameters + Adam state to a raw file and build/pth_writer (libtorch, host/pth_writer.cpp) turns it into the two torch::save
// archives in the background; older *_latest_* files are removed like the reference's clean-up loop.  Without the helper (no
// libtorch at build time) only the engine's own .bin checkpoints are written, with one notice.
struct PthSaver {
  std::string helper;
  bool warned = false;
  // One background worker and at most one waiting request: a "latest" save that arrives while the previous conversion is still
  // running replaces the waiting one (iterations of a few milliseconds would otherwise each wait ~0.4 s for a libtorch process to
  // start), so the newest state is always the one that gets written; the final save drains the queue before it returns.
  struct Req { std::string raw, cmd, model_name, opt_name; std::filesystem::path folder; bool rotate; };
  std::thread worker;
  std::mutex mu;
  std::condition_variable cv;
  std::optional<Req> pending;
  bool running = false, stop = false;
  explicit PthSaver(const char *argv0) {
    std::error_code ec;
    auto dir = std::filesystem::weakly_canonical(std::filesystem::path(argv0), ec).parent_path();
    if (std::filesystem::exists(dir / "pth_writer")) helper = (dir / "pth_writer").string();
  }
  ~PthSaver() {
    wait();
    { std::lock_guard<std::mutex> l(mu); stop = true; }
    cv.notify_all();
    if (worker.joinable()) worker.join();
  }
  void wait() {
    std::unique_lock<std::mutex> l(mu);
    cv.wait(l, [&] { return !pending && !running; });
  }
  void run(const Req &r) {
    if (std::system(r.cmd.c_str()) != 0) std::cerr << "pth_writer failed: " << r.cmd << std::endl;
    std::remove(r.raw.c_str());
    if (!r.rotate) return;
    std::error_code ec;   // :549-563
    for (const auto &e : std::filesystem::directory_iterator(r.folder, ec)) {
      const std::string fn = e.path().filename().string();
      const bool old_model = fn.rfind("model_latest_", 0) == 0 && fn.size() > 4 && fn.substr(fn.size() - 4) == ".pth" && fn != r.model_name;
      const bool old_opt = fn.rfind("optimizer_latest_", 0) == 0 && fn.size() > 4 && fn.substr(fn.size() - 4) == ".pth" && fn != r.opt_name;
      if (old_model || old_opt) std::filesystem::remove(e.path(), ec);
    }
  }
  void loop() {
    std::unique_lock<std::mutex> l(mu);
    for (;;) {
      cv.wait(l, [&] { return pending || stop; });
      if (!pending) return;
      const Req r = *pending;
      pending.reset();
      running = true;
      l.unlock();
      run(r);
      l.lock();
      running = false;
      cv.notify_all();
    }
  }
  // kind 0: ppo agent, 1: ac_ppo agent (mean / std / action range are part of that module's state)
  void save(const std::filesystem::path &folder, const std::string &model_name, const std::string &opt_name, int kind, int O, int A, long long step,
            float lr, float eps, const std::vector<float> &params, const std::vector<float> &m, const std::vector<float> &v, float act_high = 0.f,
            float act_low = 0.f, const std::vector<float> *mean = nullptr, const std::vector<float> *sd = nullptr, bool rotate = true) {
    if (helper.empty()) {
      if (!warned) std::cerr << "pth_writer not built (no libtorch found at build time): .pth files are not written" << std::endl;
      warned = true;
      return;
    }
    Req r;
    r.raw = (folder / (model_name + ".raw")).string(); r.model_name = model_name; r.opt_name = opt_name; r.folder = folder; r.rotate = rotate;
    {
      std::ofstream f(r.raw, std::ios::binary);
      const int32_t magic = 0x50425257, k = kind, o = O, a = A, P = (int32_t)params.size();
      const int64_t st = step;
      f.write((const char *)&magic, 4); f.write((const char *)&k, 4); f.write((const char *)&o, 4); f.write((const char *)&a, 4); f.write((const char *)&P, 4);
      f.write((const char *)&st, 8); f.write((const char *)&lr, 4); f.write((const char *)&eps, 4); f.write((const char *)&act_high, 4); f.write((const char *)&act_low, 4);
      f.write((const char *)params.data(), 4l * P); f.write((const char *)m.data(), 4l * P); f.write((const char *)v.data(), 4l * P);
      if (kind == 1 && mean && sd) { f.write((const char *)mean->data(), 4l * O); f.write((const char *)sd->data(), 4l * O); }
    }
    r.cmd = "'" + helper + "' write '" + r.raw + "' '" + (folder / model_name).string() + "' '" + (folder / opt_name).string() + "'";
    if (!rotate) wait();   // the final files do not supersede the last "latest" ones: let those be written first
    {
      std::lock_guard<std::mutex> l(mu);
      if (!worker.joinable()) worker = std::thread([this] { loop(); });
      if (pending) std::remove(pending->raw.c_str());   // superseded before its conversion started
      pending = std::move(r);
    }
    cv.notify_all();
    if (!rotate) wait();
  }
};

// Counter shared by the ranks of one node: what include/tcp_store.h's ZeroMQ server / clients provide in the reference
// (reset by rank 0 at the start of an iteration, incremented by every collector that has finished, read by the collectors
// that are still running, src/ac_ppo_continuous_action.cpp:629-631,680-693).  An 8-byte file mapped by every rank.
struct SharedCounter {
  long long *p = nullptr;
  int fd = -1;
  bool open(const std::string &path, bool create) {
    fd = ::open(path.c_str(), create ? (O_RDWR | O_CREAT) : O_RDWR, 0600);
    if (fd < 0) return false;
    if (create && ftruncate(fd, 8) != 0) return false;
    void *m = mmap(nullptr, 8, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (m == MAP_FAILED) return false;
    p = (long long *)m;
    return true;
  }
  void reset() { __atomic_store_n(p, 0ll, __ATOMIC_SEQ_CST); }
  void add(long long n) { __atomic_fetch_add(p, n, __ATOMIC_SEQ_CST); }
  long long get() const { return __atomic_load_n(p, __ATOMIC_SEQ_CST); }
  ~SharedCounter() { if (p) munmap(p, 8); if (fd >= 0) ::close(fd); }
};

}  // namespace cli
