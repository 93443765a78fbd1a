// main.cpp — command line of the renderer.  Same flags, defaults, clamps and
// messages as the reference's main.cpp:29-165:
//   -s N  -ns N  -dx Nx  -dy Ny  -h  -v  -g      PPM (P3) on stdout, timing on stderr.
// Documented normalisations: Q1 `-ns N` really takes N samples per
// pixel (the reference stores N and traces 1); Q2 `-s` accepts 0..4 (the reference's
// range test only lets 4 through).  Additive flags: -o FILE, -d DEVICE, -gpus N, -a ASSET_DIR, -j (JSON stats).
#include <chrono>
#include <cstdio>
#include <iostream>
#include <stdexcept>
#include <string>

#include "director.h"
#include "input_parser.h"

#define Nx_MIN (320)
#define Ny_MIN (200)
#define Nx_MAX (3840)
#define Ny_MAX (2240)
#define Nscene_MAX (4)
#define Ns_MAX (1024 * 10)

static bool parse_int(const std::string& s, int& out) {
    try {
        std::size_t pos;
        out = std::stoi(s, &pos);
        return true;
    } catch (std::invalid_argument const&) {
        return false;
    } catch (std::out_of_range const&) {
        return false;
    }
}

int main(int argc, char* argv[]) {
    int Nx = 1200, Ny = 600, Nscene = 4, Ns = 20;   // main.cpp:34-37
    bool Qverbose = false, Qdebug = false;
    InputParser cl_input(argc, argv);
    if (cl_input.cmdOptionExists("-h") || cl_input.cmdOptionExists("--help")) {
        std::cerr << std::endl << " HELP - " << argv[0] << std::endl;
        std::cerr << R"(
    -s N           Scene Selection number N (N: 0, 1, 2, etc.)
    -ns N          Sample each pixel N times (N: 1, 2, etc.)
    -dx Nx         Output image width (x dimension)
    -dy Ny         Output image height (y dimension)

    -h             This help message.
    -v             Verbose output.
    -g             Debug output.

    -o FILE        Write the PPM to FILE instead of stdout.
    -d N           CUDA device index (default 0).
    -gpus N        Render on N GPUs of this box (devices d .. d+N-1): samples or image tiles are sharded,
                   one NCCL reduce / gather per frame.
    -denoise N     Pass the frame through N (1..8) passes of the a-trous stand-in for the reference's OptiX AI denoiser.
    -a DIR         Asset directory holding earthmap.ppm (default "assets").
    -j             Print one JSON line with render statistics on stderr.
    -f FILE        Render the scene stored in FILE (scene wire format) instead of scene -s N.
    -w FILE        Write the scene in use to FILE (scene wire format).

)";
        return EXIT_SUCCESS;
    }
    if (cl_input.cmdOptionExists("-v")) Qverbose = true;
    if (cl_input.cmdOptionExists("-g")) Qdebug = true;

    const std::string& sceneNumber = cl_input.getCmdOption("-s");
    if (!sceneNumber.empty()) {
        int x;
        if (!parse_int(sceneNumber, x)) {
            std::cerr << "Invalid scene number: " << sceneNumber << std::endl;
        } else if (x >= 0 && x <= Nscene_MAX) {   // Q2
            Nscene = x;
        } else {
            std::cerr << "WARNING: Scene number " << x << " out of range. Maximum scene number: " << Nscene_MAX << std::endl;
            std::cerr << "WARNING: Using a scene value of " << Nscene << std::endl;
        }
    }
    const std::string& numberOfSamples = cl_input.getCmdOption("-ns");
    if (!numberOfSamples.empty()) {
        int x;
        if (!parse_int(numberOfSamples, x)) {
            std::cerr << "Invalid number of samples: " << numberOfSamples << std::endl;
        } else if ((x > 0) && (x <= Ns_MAX)) {
            Ns = x;
        } else {
            std::cerr << "WARNING: Number of samples " << x << " is out of range. ";
            if (x > Ns_MAX) Ns = Ns_MAX;
            std::cerr << "Using a value of " << Ns << std::endl;
        }
    }
    const std::string& dimWidth = cl_input.getCmdOption("-dx");
    if (!dimWidth.empty()) {
        int x;
        if (!parse_int(dimWidth, x)) {
            std::cerr << "Invalid image width (-dx): " << dimWidth << std::endl;
        } else if (x >= Nx_MIN && x <= Nx_MAX) {
            Nx = x;
        } else {
            std::cerr << "WARNING: Width (-dx) " << x << " out of range. ";
            if (x > Nx_MAX) Nx = Nx_MAX;
            if (x < Nx_MIN) Nx = Nx_MIN;
            std::cerr << "Using a value of " << Nx << std::endl;
        }
    }
    const std::string& dimHeight = cl_input.getCmdOption("-dy");
    if (!dimHeight.empty()) {
        int x;
        if (!parse_int(dimHeight, x)) {
            std::cerr << "Invalid image height (-dy): " << dimHeight << std::endl;
        } else if (x >= Ny_MIN && x <= Ny_MAX) {
            Ny = x;
        } else {
            std::cerr << "WARNING: Width (-dy) " << x << " out of range. ";   // (sic) main.cpp:131
            if (x > Ny_MAX) Ny = Ny_MAX;
            if (x < Ny_MIN) Ny = Ny_MIN;
            std::cerr << "Using a value of " << Ny << std::endl;
        }
    }
    int device = 0;
    const std::string& dev = cl_input.getCmdOption("-d");
    if (!dev.empty()) parse_int(dev, device);
    int gpus = 1;
    const std::string& ngpu = cl_input.getCmdOption("-gpus");
    if (!ngpu.empty()) parse_int(ngpu, gpus);
    const std::string& out = cl_input.getCmdOption("-o");
    const std::string& assets = cl_input.getCmdOption("-a");
    const std::string& scene_in = cl_input.getCmdOption("-f");
    const std::string& scene_out = cl_input.getCmdOption("-w");

    Director director(Qverbose, Qdebug);
    director.setDevice(device);
    director.setGpuCount(gpus);
    if (!assets.empty()) director.setAssetDir(assets);

    auto start = std::chrono::system_clock::now();
    director.init(Nx, Ny, Ns);
    if (Qverbose) {
        std::cerr << "INFO: Output image dimensions: " << Nx << 'x' << Ny << std::endl;
        std::cerr << "INFO: Number of rays sent per pixel: " << Ns << std::endl;
        std::cerr << "INFO: Scene number selected: " << Nscene << std::endl;
    }
    if (!scene_in.empty()) director.createSceneFromFile(scene_in.c_str());
    else director.createScene(Nscene);
    if (!scene_out.empty()) director.saveScene(scene_out.c_str());
    director.renderFrame();
    int dn = 0;
    const std::string& dns = cl_input.getCmdOption("-denoise");
    if (!dns.empty() && parse_int(dns, dn) && dn > 0) director.denoise(dn > 8 ? 8 : dn);
    auto stop = std::chrono::system_clock::now();
    auto time_seconds = std::chrono::duration<float>(stop - start).count();
    std::cerr << "INFO: Took " << time_seconds << " seconds." << std::endl;   // main.cpp:160 (Q35: whole run)
    const rtw_stats& st = director.stats();
    if (Qverbose)
        std::cerr << "INFO: render " << st.render_ms << " ms, LBVH build " << st.build_ms << " ms, "
                  << (double)st.paths / (st.render_ms * 1e3) << " Mpaths/s, "
                  << (double)(st.rays_closest + st.rays_shadow) / (st.render_ms * 1e3) << " Mrays/s" << std::endl;
    if (cl_input.cmdOptionExists("-j"))
        fprintf(stderr,
                "{\"scene\": %d, \"Nx\": %d, \"Ny\": %d, \"spp\": %d, \"paths\": %llu, \"rays_closest\": %llu, "
                "\"rays_shadow\": %llu, \"render_ms\": %.3f, \"build_ms\": %.3f, \"kernel_launches\": %llu, \"iterations\": %llu}\n",
                Nscene, Nx, Ny, Ns, (unsigned long long)st.paths, (unsigned long long)st.rays_closest,
                (unsigned long long)st.rays_shadow, st.render_ms, st.build_ms, (unsigned long long)st.kernel_launches,
                (unsigned long long)st.iterations);
    director.printPPM(out.empty() ? "-" : out.c_str());
    director.destroy();
    return EXIT_SUCCESS;
}
