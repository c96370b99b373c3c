// Oracle build only (TEST INFRASTRUCTURE): stands in for htslib's BGZF so that rpvg's threaded_output_writer.cpp compiles
// unmodified without htslib. Mode "wg" (gzip) goes through zlib, "wu" (uncompressed) through stdio: the bytes rpvg's
// writers hand over arrive in the file unchanged (bgzf is a gzip stream too, only blocked differently).
#pragma once
#include <cstdio>
#include <cstring>
#include <zlib.h>

struct BGZF {
    gzFile gz = nullptr;
    FILE * plain = nullptr;
};

inline BGZF * bgzf_open(const char * path, const char * mode) {
    BGZF * fp = new BGZF();
    if (std::strchr(mode, 'u')) { fp->plain = std::fopen(path, "wb"); } else { fp->gz = gzopen(path, "wb"); }
    if (!fp->plain && !fp->gz) { delete fp; return nullptr; }
    return fp;
}

inline long bgzf_write(BGZF * fp, const void * data, size_t length) {
    if (length == 0) { return 0; }
    if (fp->plain) { return std::fwrite(data, 1, length, fp->plain) == length ? static_cast<long>(length) : -1; }
    return gzwrite(fp->gz, data, static_cast<unsigned>(length)) > 0 ? static_cast<long>(length) : -1;
}

inline int bgzf_close(BGZF * fp) {
    int status = fp->plain ? std::fclose(fp->plain) : (gzclose(fp->gz) == Z_OK ? 0 : -1);
    delete fp;
    return status;
}
