// director.cpp — see director.h.  Errors: the reference throws sutil::Exception that
// nothing catches (→ std::terminate) and exits on an unknown scene (Director.cpp:955-958);
// here every C-ABI status is checked and reported before exit(EXIT_FAILURE).
#include "director.h"

#include <cstdio>
#include <cstdlib>
#include <iostream>

void Director::check(int rc, const char* what) {
    if (rc == RTW_OK) return;
    std::cerr << "ERROR: " << what << " failed (" << rc << "): " << (m_multi ? rtw_multi_last_error(m_multi) : rtw_last_error(m_ctx)) << std::endl;
    std::exit(EXIT_FAILURE);
}

void Director::init(unsigned int width, unsigned int height, unsigned int samples) {
    m_Nx = (int)width;
    m_Ny = (int)height;
    m_Ns = (int)samples;
    if (m_gpus > 1) {
        std::vector<int> devices;
        for (int g = 0; g < m_gpus; g++) devices.push_back(m_device + g);
        check(rtw_create_multi(devices.data(), m_gpus, &m_multi), "rtw_create_multi");
        if (_debug) check(rtw_multi_set_option(m_multi, "profile", 1), "rtw_multi_set_option(profile)");
        return;
    }
    check(rtw_create(m_device, &m_ctx), "rtw_create");
    if (_debug) check(rtw_set_option(m_ctx, "profile", 1), "rtw_set_option(profile)");
}

void Director::createScene(unsigned int sceneNumber) {
    int rc = rtw_scene_create((int)sceneNumber, m_Nx, m_Ny, m_assets.c_str(), 0, &m_scene);
    if (rc != RTW_OK) std::exit(EXIT_FAILURE);  // "ERROR: Scene N unknown." already printed
    uploadScene();
    if (_verbose) std::cerr << "INFO: Scene description: " << rtw_scene_description(m_scene) << std::endl;  // Director.cpp:965-968
}

void Director::uploadScene() {
    if (m_multi) {
        check(rtw_multi_upload_scene(m_multi, rtw_scene_get_desc(m_scene)), "rtw_multi_upload_scene");
        check(rtw_multi_build_accel(m_multi), "rtw_multi_build_accel");
    } else {
        check(rtw_upload_scene(m_ctx, rtw_scene_get_desc(m_scene)), "rtw_upload_scene");
        check(rtw_build_accel(m_ctx), "rtw_build_accel");
    }
}

void Director::createSceneFromFile(const char* path) {
    int rc = rtw_scene_load(path, &m_scene);
    if (rc != RTW_OK) {
        std::cerr << "ERROR: cannot load scene " << path << ": " << rtw_scene_io_error() << std::endl;
        std::exit(EXIT_FAILURE);
    }
    uploadScene();
    if (_verbose) std::cerr << "INFO: Scene description: " << rtw_scene_description(m_scene) << std::endl;
}

void Director::saveScene(const char* path) {
    int rc = rtw_scene_save(rtw_scene_get_desc(m_scene), rtw_scene_description(m_scene), path);
    if (rc != RTW_OK) {
        std::cerr << "ERROR: cannot save scene " << path << ": " << rtw_scene_io_error() << std::endl;
        std::exit(EXIT_FAILURE);
    }
}

void Director::renderFrame() {
    m_frame.assign((size_t)m_Nx * m_Ny * 4, 0.f);
    if (m_multi) {
        check(rtw_multi_render(m_multi, m_Nx, m_Ny, m_Ns, m_maxRayDepth, m_frame.data()), "rtw_multi_render");
        check(rtw_multi_get_stats(m_multi, &m_stats, nullptr, nullptr, nullptr), "rtw_multi_get_stats");
        return;
    }
    check(rtw_render(m_ctx, m_Nx, m_Ny, m_Ns, m_maxRayDepth, m_frame.data()), "rtw_render");
    check(rtw_get_stats(m_ctx, &m_stats), "rtw_get_stats");
}

void Director::denoise(int iterations, float sigma) {
    if (!m_ctx) { std::cerr << "WARNING: -denoise needs a one-GPU context; frame left as rendered" << std::endl; return; }
    check(rtw_denoise(m_ctx, m_Nx, m_Ny, iterations, sigma, m_frame.data(), m_frame.data()), "rtw_denoise");
}

void Director::printPPM(const char* path) {
    int rc = rtw_write_ppm(path, m_Nx, m_Ny, m_frame.data());
    if (rc != RTW_OK) {
        std::cerr << "ERROR: cannot write " << path << std::endl;
        std::exit(EXIT_FAILURE);
    }
}

void Director::destroy() {
    if (m_scene) { rtw_scene_free(m_scene); m_scene = nullptr; }
    if (m_ctx) { rtw_destroy(m_ctx); m_ctx = nullptr; }
    if (m_multi) { rtw_multi_destroy(m_multi); m_multi = nullptr; }
}
