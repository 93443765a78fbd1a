// director.h — C++ host mirror of the reference's Director (Director.h:118-126)
// on top of the C-ABI (include/rtw.h).  Same call order as main.cpp:145-164:
// init → createScene → renderFrame → printPPM → destroy.
#pragma once
#include <string>
#include <vector>
#include "../../../include/rtw.h"

class Director {
public:
    Director(bool verbose = false, bool debug = false) : _verbose(verbose), _debug(debug) {}
    ~Director() { destroy(); }

    void init(unsigned int width, unsigned int height, unsigned int samples);   // Director.cpp:33
    void createScene(unsigned int sceneNumber);                                  // Director.cpp:951
    void createSceneFromFile(const char* path);    // scene wire format (rtw_scene_load) instead of a built-in scene function
    void saveScene(const char* path);              // rtw_scene_save of the scene in use
    void renderFrame();                                                          // Director.cpp:971
    void denoise(int iterations, float sigma = 0.25f);   // stand-in for Director.cpp:988-1000 (a-trous filter, rtw_denoise); 1-GPU contexts
    void printPPM(const char* path = "-");                                       // Director.cpp:1010
    void destroy();                                                              // Director.cpp:86

    void setDevice(int device) { m_device = device; }
    void setGpuCount(int n) { m_gpus = n < 1 ? 1 : n; }   // > 1: rtw_create_multi over devices device .. device + n - 1
    void setAssetDir(const std::string& d) { m_assets = d; }
    const rtw_stats& stats() const { return m_stats; }
    const std::vector<float>& frame() const { return m_frame; }

private:
    void check(int rc, const char* what);
    void uploadScene();
    bool _verbose, _debug;
    int m_device = 0;
    int m_Nx = 0, m_Ny = 0, m_Ns = 0;
    int m_maxRayDepth = 20;   // Director.cpp:42 (Q9)
    std::string m_assets = "assets";
    int m_gpus = 1;
    rtw_ctx* m_ctx = nullptr;
    rtw_multi* m_multi = nullptr;
    rtw_scene* m_scene = nullptr;
    rtw_stats m_stats{};
    std::vector<float> m_frame;
};
