// See probs_io.hpp. Format restated from rpvg:src/threaded_output_writer.cpp:42-95.
#include "probs_io.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <utility>

#include <zlib.h>

namespace rpvg_b200_host {

namespace {

// the reference's out_precision_digits (rpvg:src/threaded_output_writer.cpp:6)
constexpr uint32_t kOutPrecisionDigits = 8;

[[noreturn]] void fail(const std::string & filename, uint64_t line, const std::string & what) {

    throw std::runtime_error(filename + ":" + std::to_string(line) + ": " + what);
}

// One line of a gzip stream, without the newline; false at end of file.
bool readLine(gzFile file, std::string * line) {

    line->clear();
    char buffer[1 << 16];

    while (gzgets(file, buffer, sizeof(buffer)) != nullptr) {

        const size_t len = std::strlen(buffer);
        line->append(buffer, len);

        if (len > 0 && buffer[len - 1] == '\n') {
            line->pop_back();
            if (!line->empty() && line->back() == '\r') { line->pop_back(); }
            return true;
        }
    }

    return !line->empty();
}

std::vector<std::string> split(const std::string & text, const char sep) {

    std::vector<std::string> parts;
    size_t begin = 0;

    while (begin <= text.size()) {
        const size_t end = text.find(sep, begin);
        if (end == std::string::npos) {
            parts.push_back(text.substr(begin));
            break;
        }
        parts.push_back(text.substr(begin, end - begin));
        begin = end + 1;
    }

    return parts;
}

double parseDouble(const std::string & text, const std::string & filename, uint64_t line) {

    char * end = nullptr;
    const double value = std::strtod(text.c_str(), &end);
    if (text.empty() || *end != '\0') { fail(filename, line, "not a number: '" + text + "'"); }
    return value;
}

uint64_t parseUnsigned(const std::string & text, const std::string & filename, uint64_t line) {

    char * end = nullptr;
    const unsigned long long value = std::strtoull(text.c_str(), &end, 10);
    if (text.empty() || *end != '\0' || text[0] == '-') { fail(filename, line, "not an unsigned integer: '" + text + "'"); }
    return value;
}

// like `stream << setprecision(digits) << value` (default float format = %g with that many significant digits)
std::string formatDouble(const double value, const uint32_t digits) {

    char buffer[64];
    std::snprintf(buffer, sizeof(buffer), "%.*g", static_cast<int>(digits), value);
    return buffer;
}

}

ProbsDump readProbsDump(const std::string & filename) {

    gzFile file = gzopen(filename.c_str(), "rb");
    if (!file) { throw std::runtime_error(filename + ": cannot open"); }

    ProbsDump dump;
    dump.read_offset.push_back(0);
    dump.path_offset.push_back(0);
    dump.row_ptr.push_back(0);

    std::string line;
    uint64_t line_number = 0;
    bool expect_paths = false;
    bool in_cluster = false;
    uint32_t cluster_paths = 0;

    std::vector<std::pair<uint32_t, double> > entries;

    try {

        while (readLine(file, &line)) {

            ++line_number;

            if (line == "#") {

                if (expect_paths) { fail(filename, line_number, "cluster header without path line"); }
                if (in_cluster) {
                    dump.read_offset.push_back(dump.noise.size());
                    dump.path_offset.push_back(dump.path_name.size());
                }
                expect_paths = true;
                in_cluster = true;
                continue;
            }

            if (!in_cluster) { fail(filename, line_number, "text before the first '#'"); }
            if (line.empty()) { fail(filename, line_number, "empty line"); }

            if (expect_paths) {

                cluster_paths = 0;

                for (auto & item: split(line, ' ')) {

                    // name,length,effective_length — the name itself may contain commas: the last two fields count
                    const size_t c2 = item.rfind(',');
                    const size_t c1 = (c2 == std::string::npos || c2 == 0) ? std::string::npos : item.rfind(',', c2 - 1);
                    if (c1 == std::string::npos) { fail(filename, line_number, "path entry is not name,length,effective_length: '" + item + "'"); }

                    dump.path_name.push_back(item.substr(0, c1));
                    dump.path_length.push_back(static_cast<uint32_t>(parseUnsigned(item.substr(c1 + 1, c2 - c1 - 1), filename, line_number)));
                    dump.path_effective_length.push_back(parseDouble(item.substr(c2 + 1), filename, line_number));
                    ++cluster_paths;
                }

                if (cluster_paths == 0) { fail(filename, line_number, "cluster without paths"); }
                expect_paths = false;
                continue;
            }

            // read class: count noise prob:ids ...
            const auto fields = split(line, ' ');
            if (fields.size() < 2) { fail(filename, line_number, "read class needs a count and a noise probability"); }

            const uint64_t count = parseUnsigned(fields[0], filename, line_number);
            if (count > 0xffffffffull) { fail(filename, line_number, "read count does not fit 32 bits"); }

            entries.clear();

            for (size_t f = 2; f < fields.size(); ++f) {

                const size_t colon = fields[f].find(':');
                if (colon == std::string::npos) { fail(filename, line_number, "expected prob:ids, got '" + fields[f] + "'"); }

                const double prob = parseDouble(fields[f].substr(0, colon), filename, line_number);

                for (auto & id_text: split(fields[f].substr(colon + 1), ',')) {
                    const uint64_t id = parseUnsigned(id_text, filename, line_number);
                    if (id >= cluster_paths) { fail(filename, line_number, "path id " + id_text + " out of range (cluster has " + std::to_string(cluster_paths) + " paths)"); }
                    entries.emplace_back(static_cast<uint32_t>(id), prob);
                }
            }

            std::sort(entries.begin(), entries.end());

            for (size_t e = 0; e < entries.size(); ++e) {
                if (e > 0 && entries[e].first == entries[e - 1].first) { fail(filename, line_number, "path id " + std::to_string(entries[e].first) + " listed twice"); }
                dump.col.push_back(entries[e].first);
                dump.prob.push_back(entries[e].second);
            }

            dump.row_ptr.push_back(dump.col.size());
            dump.noise.push_back(parseDouble(fields[1], filename, line_number));
            dump.count.push_back(static_cast<uint32_t>(count));
        }

        if (expect_paths) { fail(filename, line_number, "file ends after a cluster header"); }

    } catch (...) {
        gzclose(file);
        throw;
    }

    gzclose(file);

    if (in_cluster) {
        dump.read_offset.push_back(dump.noise.size());
        dump.path_offset.push_back(dump.path_name.size());
    }

    return dump;
}

void writeProbsDump(const std::string & filename, const ProbsDump & dump, const double prob_precision) {

    const uint32_t prob_digits = std::max(kOutPrecisionDigits, static_cast<uint32_t>(std::ceil(-1 * std::log10(prob_precision))));

    gzFile file = gzopen(filename.c_str(), "wb");
    if (!file) { throw std::runtime_error(filename + ": cannot open for writing"); }

    std::string text;
    std::vector<std::pair<double, uint32_t> > entries;

    for (uint32_t c = 0; c < dump.numClusters(); ++c) {

        const uint64_t r0 = dump.read_offset[c], r1 = dump.read_offset[c + 1];
        const uint64_t p0 = dump.path_offset[c], p1 = dump.path_offset[c + 1];

        if (r0 == r1) { continue; }   // the reference skips clusters without read classes

        text = "#\n";

        for (uint64_t p = p0; p < p1; ++p) {
            if (p > p0) { text += ' '; }
            text += dump.path_name[p] + "," + std::to_string(dump.path_length[p]) + "," + formatDouble(dump.path_effective_length[p], kOutPrecisionDigits);
        }
        text += '\n';

        for (uint64_t r = r0; r < r1; ++r) {

            text += std::to_string(dump.count[r]) + " " + formatDouble(dump.noise[r], prob_digits);

            // groups of equal probability, ascending by (probability, ids) like ReadPathProbabilities::pathProbs()
            entries.clear();
            for (uint64_t e = dump.row_ptr[r]; e < dump.row_ptr[r + 1]; ++e) { entries.emplace_back(dump.prob[e], dump.col[e]); }
            std::sort(entries.begin(), entries.end());

            for (size_t e = 0; e < entries.size(); ++e) {
                if (e == 0 || entries[e].first != entries[e - 1].first) {
                    text += " " + formatDouble(entries[e].first, prob_digits) + ":" + std::to_string(entries[e].second);
                } else {
                    text += "," + std::to_string(entries[e].second);
                }
            }
            text += '\n';
        }

        if (gzwrite(file, text.data(), static_cast<unsigned>(text.size())) != static_cast<int>(text.size())) {
            gzclose(file);
            throw std::runtime_error(filename + ": write failed");
        }
    }

    if (gzclose(file) != Z_OK) { throw std::runtime_error(filename + ": close failed"); }
}

}

namespace {

struct ProbsHandle {

    rpvg_b200_host::ProbsDump dump;
    std::string names;
};

void setError(char * error, uint64_t error_len, const std::string & message) {

    if (error && error_len > 0) {
        std::strncpy(error, message.c_str(), error_len - 1);
        error[error_len - 1] = '\0';
    }
}

}

extern "C" void * rpvg_b200_probs_read(const char * filename, rpvg_b200_probs_view * view, char * error, uint64_t error_len) {

    if (!filename || !view) {
        setError(error, error_len, "NULL argument");
        return nullptr;
    }

    try {

        auto * handle = new ProbsHandle();
        handle->dump = rpvg_b200_host::readProbsDump(filename);

        auto & d = handle->dump;
        for (size_t p = 0; p < d.path_name.size(); ++p) {
            if (p > 0) { handle->names += '\n'; }
            handle->names += d.path_name[p];
        }

        view->n_clusters = d.numClusters();
        view->n_reads = d.noise.size();
        view->n_paths = d.path_name.size();
        view->nnz = d.col.size();
        view->read_offset = d.read_offset.data();
        view->path_offset = d.path_offset.data();
        view->row_ptr = d.row_ptr.data();
        view->col = d.col.data();
        view->prob = d.prob.data();
        view->noise = d.noise.data();
        view->count = d.count.data();
        view->path_length = d.path_length.data();
        view->path_effective_length = d.path_effective_length.data();
        view->path_names = handle->names.c_str();

        return handle;

    } catch (const std::exception & e) {
        setError(error, error_len, e.what());
        return nullptr;
    }
}

extern "C" void rpvg_b200_probs_free(void * handle) {

    delete static_cast<ProbsHandle *>(handle);
}

extern "C" int rpvg_b200_probs_write(const char * filename, const rpvg_b200_probs_view * view, double prob_precision, char * error, uint64_t error_len) {

    if (!filename || !view) {
        setError(error, error_len, "NULL argument");
        return -1;
    }

    try {

        rpvg_b200_host::ProbsDump d;
        d.read_offset.assign(view->read_offset, view->read_offset + view->n_clusters + 1);
        d.path_offset.assign(view->path_offset, view->path_offset + view->n_clusters + 1);
        d.row_ptr.assign(view->row_ptr, view->row_ptr + view->n_reads + 1);
        d.col.assign(view->col, view->col + view->nnz);
        d.prob.assign(view->prob, view->prob + view->nnz);
        d.noise.assign(view->noise, view->noise + view->n_reads);
        d.count.assign(view->count, view->count + view->n_reads);
        d.path_length.assign(view->path_length, view->path_length + view->n_paths);
        d.path_effective_length.assign(view->path_effective_length, view->path_effective_length + view->n_paths);

        const std::string names = view->path_names ? view->path_names : "";
        size_t begin = 0;
        for (uint64_t p = 0; p < view->n_paths; ++p) {
            const size_t end = names.find('\n', begin);
            d.path_name.push_back(names.substr(begin, end == std::string::npos ? std::string::npos : end - begin));
            begin = end == std::string::npos ? names.size() : end + 1;
        }

        rpvg_b200_host::writeProbsDump(filename, d, prob_precision);
        return 0;

    } catch (const std::exception & e) {
        setError(error, error_len, e.what());
        return -1;
    }
}
