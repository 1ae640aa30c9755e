// tfevents_selftest — writes a fixed set of events with include/tfevents_writer.h; tests/test_tfevents.py reads the file
// back with TensorBoard's own loader.  usage: tfevents_selftest <path containing "tfevents">
#include <iostream>

#include "tfevents_writer.h"

int main(int argc, char **argv) {
  if (argc < 2) { std::cerr << "usage: tfevents_selftest <file>" << std::endl; return 2; }
  try {
    tfevents::Writer w(argv[1]);
    w.add_text("hyperparameters", 0, "env_id: Humanoid-v4\nnum_envs: 8192\n");
    for (int i = 1; i <= 5; i++) {
      w.add_scalar("charts/SPS", (int64_t)i * 262144, 1.5e6f + (float)i);
      w.add_scalar("losses/value_loss", (int64_t)i * 262144, 0.25 / i);
    }
    w.add_scalar("charts/episodic_return", 3000000000LL, -12.5f);   // step beyond 32 bits
    w.flush();
  } catch (const std::exception &e) {
    std::cerr << e.what() << std::endl;
    return 1;
  }
  return 0;
}
