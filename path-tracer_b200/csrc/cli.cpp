// cli.cpp — `path-tracer` command line.  Mirrors the reference's clap surface (src/config/mod.rs:14-52) and
// run_render (src/main.rs:32-52): `path-tracer render <input> [-o FILE|$OUTPUT] [-q] [-v] [--debug-textures]
// [-p FILE|$PROFILE]`, errors printed to stderr with exit code 2 (main.rs:14-21).  Additive flags: --seed,
// --stats, --snapshot-every, --gpus N|all (one frame split by sample index over N GPUs of the box through
// pt_render_multi; also $PT_B200_GPUS).  The SFML viewer (renderer/viewer.rs) needs a window + GL: unavailable on this box; -v
// delivers the same feed headless instead: the running mean image is written over <OUTPUT> after every
// --snapshot-every samples per pixel (default 16).  --debug-textures writes the seven AOV PNGs of
// renderer/debug_renderer.rs into the working directory.  `convert <INPUT> <OUTPUT>` (glTF -> isf, scene/gltf.rs) is
// host code (gltf_convert.cpp); `convert --binary` also writes <OUTPUT>/scene.isfb, and `pack <INPUT> <OUTPUT.isfb>`
// turns any isf / isfb scene into the binary side-car (raw arrays + the host SAH tree unless --no-accel;
// isf_binary.cpp), which `render` accepts wherever it accepts scene.isf.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pt_b200.h"

static int usage(const char *msg) {
  if (msg) std::fprintf(stderr, "error: %s\n\n", msg);
  std::fprintf(stderr,
               "Path-trace awesome things\n\nUsage: path-tracer <COMMAND>\n\nCommands:\n"
               "  render   <INPUT> [-o, --output <OUTPUT>] [-q, --quiet] [-v, --viewer] [--debug-textures]\n"
               "           [-p, --profile <PROFILE>] [--seed <N>] [--stats <FILE>] [--snapshot-every <SPP>]\n"
               "           [--gpus <N|all>]\n"
               "  convert  <INPUT> <OUTPUT> [--binary]   glTF 2.0 (.gltf / .glb) -> <OUTPUT>/scene.isf + texture PNGs\n"
               "           (--binary: also <OUTPUT>/scene.isfb)\n"
               "  pack     <INPUT> <OUTPUT.isfb> [--no-accel]   isf scene -> binary side-car with its BVH\n");
  return 2;
}

int main(int argc, char **argv) {
  if (argc < 2) return usage("a subcommand is required");
  std::string cmd = argv[1];
  if (cmd == "-h" || cmd == "--help") {
    usage(nullptr);
    return 0;
  }
  if (cmd == "convert" || cmd == "pack") {  // config/mod.rs:44-52 + main.rs:54-57
    std::vector<std::string> pos;
    bool binary = false, accel = true;
    for (int i = 2; i < argc; ++i) {
      std::string a = argv[i];
      if (a == "-h" || a == "--help") {
        usage(nullptr);
        return 0;
      }
      if (cmd == "convert" && a == "--binary") { binary = true; continue; }
      if (cmd == "pack" && a == "--no-accel") { accel = false; continue; }
      if (!a.empty() && a[0] == '-') return usage(("unexpected argument '" + a + "'").c_str());
      pos.push_back(a);
    }
    if (pos.size() < 2) return usage("the following required arguments were not provided: <INPUT> <OUTPUT>");
    if (pos.size() > 2) return usage(("unexpected argument '" + pos[2] + "'").c_str());
    auto fail = [] {
      std::fprintf(stderr, "%s\n", pt_last_error());
      return 2;
    };
    std::string isf = pos[0], isfb = pos[1];
    if (cmd == "convert") {
      if (pt_convert_gltf(pos[0].c_str(), pos[1].c_str()) != PT_OK) return fail();
      if (!binary) return 0;
      isf = pos[1] + "/scene.isf";
      isfb = pos[1] + "/scene.isfb";
    }
    pt_scene *host_scene = nullptr;  // device -1: parsed and finalised on the host, nothing uploaded
    if (pt_scene_load_isf(isf.c_str(), -1, &host_scene) != PT_OK) return fail();
    int rc = pt_scene_save_binary(host_scene, isfb.c_str(), accel ? 1 : 0);
    pt_scene_destroy(host_scene);
    return rc == PT_OK ? 0 : fail();
  }
  if (cmd != "render") return usage(("unrecognized subcommand '" + cmd + "'").c_str());

  const char *env_out = std::getenv("OUTPUT");
  const char *env_prof = std::getenv("PROFILE");
  std::string input, output = env_out ? env_out : "render.png", profile_path = env_prof ? env_prof : "", stats_path;
  bool quiet = false, viewer = false, debug_textures = false;
  unsigned long long seed = 0;
  unsigned snapshot_every = 16;
  std::string gpus_arg = std::getenv("PT_B200_GPUS") ? std::getenv("PT_B200_GPUS") : "1";
  for (int i = 2; i < argc; ++i) {
    std::string a = argv[i];
    auto need = [&](const char *name) -> const char * {
      if (i + 1 >= argc) {
        usage((std::string("a value is required for '") + name + "'").c_str());
        std::exit(2);
      }
      return argv[++i];
    };
    if (a == "-o" || a == "--output") output = need("--output");
    else if (a.rfind("--output=", 0) == 0) output = a.substr(9);
    else if (a == "-p" || a == "--profile") profile_path = need("--profile");
    else if (a.rfind("--profile=", 0) == 0) profile_path = a.substr(10);
    else if (a == "-q" || a == "--quiet") quiet = true;
    else if (a == "-v" || a == "--viewer") viewer = true;
    else if (a == "--debug-textures") debug_textures = true;
    else if (a == "--seed") seed = std::strtoull(need("--seed"), nullptr, 0);
    else if (a == "--stats") stats_path = need("--stats");
    else if (a == "--gpus") gpus_arg = need("--gpus");
    else if (a == "--snapshot-every") snapshot_every = (unsigned)std::strtoul(need("--snapshot-every"), nullptr, 0);
    else if (a == "-h" || a == "--help") {
      usage(nullptr);
      return 0;
    } else if (!a.empty() && a[0] == '-') return usage(("unexpected argument '" + a + "'").c_str());
    else if (input.empty()) input = a;
    else return usage(("unexpected argument '" + a + "'").c_str());
  }
  if (input.empty()) return usage("the following required arguments were not provided: <INPUT>");

  pt_profile prof;
  pt_profile_default(&prof);
  if (!profile_path.empty() && pt_profile_load_yaml(profile_path.c_str(), &prof) != PT_OK) {
    std::fprintf(stderr, "%s\n", pt_last_error());
    return 2;
  }
  if (viewer && !quiet && !debug_textures)
    std::fprintf(stderr, "note: --viewer needs a window system (unavailable on this box); writing the running image to %s every %u spp\n",
                 output.c_str(), snapshot_every);

  auto t0 = std::chrono::steady_clock::now();
  pt_scene *scene = nullptr;
  if (pt_scene_load_isf(input.c_str(), 0, &scene) != PT_OK) {
    std::fprintf(stderr, "%s\n", pt_last_error());
    return 2;
  }
  auto t1 = std::chrono::steady_clock::now();
  auto t_call = t1;
  if (debug_textures) {  // main.rs:41-44 + debug_renderer.rs:11-62: seven "<buffer_name>.png" in the working directory
    const size_t plane = (size_t)prof.width * prof.height * 3;
    std::vector<uint8_t> planes(plane * PT_AOV_COUNT);
    uint64_t n_hit = 0;
    int rc = pt_debug_textures(scene, prof.width, prof.height, planes.data(), &n_hit);
    pt_scene_destroy(scene);
    if (rc != PT_OK) {
      std::fprintf(stderr, "%s\n", pt_last_error());
      return 2;
    }
    // the reference creates a buffer on the first pixel that hits something: no hit at all, no file
    for (int k = 0; n_hit && k < PT_AOV_COUNT; ++k) {
      std::string name = std::string(pt_aov_name(k)) + ".png";
      if (pt_write_png(name.c_str(), planes.data() + plane * k, prof.width, prof.height) != PT_OK) {
        std::fprintf(stderr, "%s\n", pt_last_error());
        return 2;
      }
    }
    return 0;
  }
  if (!quiet) std::fprintf(stderr, "Rendering: %ux%u, %u samples, %u bounces\n", prof.width, prof.height, prof.samples, prof.bounces);
  std::vector<uint8_t> rgb((size_t)prof.width * prof.height * 3);
  pt_stats st{};
  struct Snap {
    const char *path;
    bool quiet;
  } snap{output.c_str(), quiet};
  auto on_snapshot = [](void *user, const uint8_t *img, uint32_t w, uint32_t h, uint32_t done) -> int {
    const Snap *s = static_cast<const Snap *>(user);
    if (pt_write_png(s->path, img, w, h) != PT_OK) return 1;  // like a closed viewer: stop sending, keep rendering
    if (!s->quiet) std::fprintf(stderr, "  %u spp -> %s\n", done, s->path);
    return 0;
  };
  // --gpus: replicas of the scene on devices 1..N-1, samples split by index (the viewer feed stays on one device)
  int n_gpus = gpus_arg == "all" ? pt_device_count() : std::atoi(gpus_arg.c_str());
  if (n_gpus < 1 || n_gpus > pt_device_count()) {
    std::fprintf(stderr, "--gpus %s: this box has %d CUDA device(s)\n", gpus_arg.c_str(), pt_device_count());
    pt_scene_destroy(scene);
    return 2;
  }
  std::vector<pt_scene *> replicas{scene};
  auto destroy_all = [&] {
    for (pt_scene *r : replicas) pt_scene_destroy(r);
  };
  int rrc = PT_OK;
  if (n_gpus > 1 && !viewer) {
    // one thread per device: a replica is a host-side repack plus one upload, independent of the others
    std::vector<pt_scene *> made(n_gpus, nullptr);
    std::vector<int> rcs(n_gpus, PT_OK);
    std::vector<std::string> errs(n_gpus);
    std::vector<std::thread> threads;
    for (int d = 1; d < n_gpus; ++d)
      threads.emplace_back([&, d] {
        rcs[d] = pt_scene_replicate(scene, d, &made[d]);
        if (rcs[d] != PT_OK) errs[d] = pt_last_error();  // thread-local message: keep it for the main thread
      });
    for (auto &t : threads) t.join();
    for (int d = 1; d < n_gpus; ++d) {
      if (made[d]) replicas.push_back(made[d]);
      if (rcs[d] != PT_OK && rrc == PT_OK) {
        rrc = rcs[d];
        std::fprintf(stderr, "device %d: %s\n", d, errs[d].c_str());
      }
    }
    if (rrc != PT_OK) {
      destroy_all();
      return 2;
    }
    std::vector<pt_stats> per(replicas.size());
    t_call = std::chrono::steady_clock::now();
    if (rrc == PT_OK) rrc = pt_render_multi(replicas.data(), (uint32_t)replicas.size(), &prof, seed, rgb.data(), nullptr, per.data());
    for (const pt_stats &p : per) {  // whole-frame counts; device time = the slowest replica
      st.samples += p.samples;
      st.extend_rays += p.extend_rays;
      st.shadow_rays += p.shadow_rays;
      st.alpha_recasts += p.alpha_recasts;
      st.alpha_tests += p.alpha_tests;
      st.kernel_launches += p.kernel_launches;
      if (p.ms_total > st.ms_total) st.ms_total = p.ms_total;
    }
  } else {
    t_call = std::chrono::steady_clock::now();
    rrc = viewer ? pt_render_progressive(scene, &prof, seed, snapshot_every, on_snapshot, &snap, rgb.data(), nullptr, &st)
                 : pt_render(scene, &prof, seed, rgb.data(), nullptr, &st);
  }
  if (rrc != PT_OK) {
    std::fprintf(stderr, "%s\n", pt_last_error());
    destroy_all();
    return 2;
  }
  auto t2 = std::chrono::steady_clock::now();
  destroy_all();
  if (pt_write_png(output.c_str(), rgb.data(), prof.width, prof.height) != PT_OK) {
    std::fprintf(stderr, "%s\n", pt_last_error());
    return 2;
  }
  double load_s = std::chrono::duration<double>(t1 - t0).count(), render_s = std::chrono::duration<double>(t2 - t1).count();
  // host wall clock around the render call itself: for several GPUs it includes the peer copies, the ordered add, the
  // tonemap and the read-back, which the per-replica device time does not
  const double call_s = std::chrono::duration<double>(t2 - t_call).count();
  if (!quiet)
    std::fprintf(stderr, "Done: %.0fs (load %.2fs, render %.3fs, %.1f Msamples/s by the host clock around the render call; %.1f by device time%s)\n",
                 load_s + render_s, load_s, render_s, call_s > 0 ? (double)st.samples / call_s / 1e6 : 0.0,
                 st.ms_total > 0 ? (double)st.samples / st.ms_total / 1e3 : 0.0, n_gpus > 1 ? " of the slowest replica" : "");
  if (!stats_path.empty()) {
    FILE *f = std::fopen(stats_path.c_str(), "w");
    if (f) {
      std::fprintf(f,
                   "{\"samples\": %llu, \"extend_rays\": %llu, \"shadow_rays\": %llu, \"alpha_tests\": %llu, "
                   "\"kernel_launches\": %llu, \"ms_render\": %.3f, \"ms_render_call_wall\": %.3f, \"load_s\": %.3f}\n",
                   (unsigned long long)st.samples, (unsigned long long)st.extend_rays, (unsigned long long)st.shadow_rays,
                   (unsigned long long)st.alpha_tests, (unsigned long long)st.kernel_launches, st.ms_total, call_s * 1e3, load_s);
      std::fclose(f);
    }
  }
  return 0;
}
