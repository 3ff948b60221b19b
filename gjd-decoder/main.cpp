/*
 * decoder — the reference's command line, on top of libgjd's C ABI.
 *
 *   ./decoder <file.jpg>
 *
 * Same contract as cpp-decoder/main.cpp:5-16 and cuda-decoder/main.cu:7-40: one positional argument,
 * exit 1 with the reference's message when it is missing, exit 0 on success, and the decoded image is
 * written as ../testing/<impl>_output_arrays/<stem>.array relative to the current directory
 * (cpp-decoder/src/parser.cpp:199-201), where <impl> is the part of this program's directory name
 * before the first '-' — which is how testing/compare.py:42-46 derives the directory it reads
 * ("gjd-decoder" -> gjd_output_arrays).  GJD_OUT_DIR overrides the directory; GJD_PPM=1 also writes
 * a P6 file beside the array.
 */
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <string>
#include <vector>

#include "../include/gjd.h"

static std::string impl_name(const char* argv0) {
    /* directory name of the executable up to the first '-' */
    char buf[4096];
    std::string exe = argv0;
    if (realpath(argv0, buf)) exe = buf;
    size_t slash = exe.find_last_of('/');
    std::string dir = slash == std::string::npos ? "." : exe.substr(0, slash);
    size_t s2 = dir.find_last_of('/');
    std::string base = s2 == std::string::npos ? dir : dir.substr(s2 + 1);
    size_t dash = base.find('-');
    if (dash != std::string::npos) base = base.substr(0, dash);
    return base.empty() ? std::string("gjd") : base;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        printf("Please provide the name of the image file to be decompressed.\n");
        return 1;
    }
    const char* path = argv[1];
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "decoder: cannot open %s: %s\n", path, strerror(errno)); return 2; }
    std::vector<uint8_t> file;
    {
        uint8_t tmp[65536];
        size_t n;
        while ((n = fread(tmp, 1, sizeof(tmp), f)) > 0) file.insert(file.end(), tmp, tmp + n);
        fclose(f);
    }
    int32_t w = 0, h = 0;
    int rc = gjd_peek(file.data(), file.size(), &w, &h);
    if (rc != GJD_OK) { fprintf(stderr, "decoder: %s: %s\n", path, gjd_strerror(rc)); return 3; }

    gjd_ctx* ctx = nullptr;
    rc = gjd_create(&ctx, 0, 0);
    if (rc != GJD_OK) { fprintf(stderr, "decoder: %s\n", gjd_strerror(rc)); return 4; }
    std::vector<uint8_t> rgb((size_t)w * h * 3);
    gjd_item it;
    memset(&it, 0, sizeof(it));
    it.jpg = file.data();
    it.len = file.size();
    it.rgb = rgb.data();
    rc = gjd_decode_batch(ctx, &it, 1);
    if (rc == GJD_OK) rc = it.status;
    if (rc != GJD_OK) {
        fprintf(stderr, "decoder: %s: %s %s\n", path, gjd_strerror(rc), gjd_last_cuda_error(ctx));
        gjd_destroy(ctx);
        return 5;
    }
    gjd_destroy(ctx);

    std::string stem = path;
    size_t slash = stem.find_last_of('/');
    if (slash != std::string::npos) stem = stem.substr(slash + 1);
    size_t dot = stem.find_last_of('.');
    if (dot != std::string::npos) stem = stem.substr(0, dot);
    std::string dir;
    if (const char* e = getenv("GJD_OUT_DIR")) dir = e;
    else dir = "../testing/" + impl_name(argv[0]) + "_output_arrays";
    mkdir(dir.c_str(), 0777);   /* the reference requires it to exist; creating it is harmless */
    std::string out = dir + "/" + stem + ".array";
    rc = gjd_write_array(out.c_str(), w, h, rgb.data());
    if (rc != GJD_OK) { fprintf(stderr, "decoder: cannot write %s\n", out.c_str()); return 6; }
    if (const char* e = getenv("GJD_PPM")) {
        if (atoi(e)) gjd_write_ppm((dir + "/" + stem + ".ppm").c_str(), w, h, rgb.data());
    }
    return 0;
}
