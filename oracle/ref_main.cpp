// ref_main.cpp -- `m17-demod-ref`: the reference's stdin loop
// (apps/m17-demod.cpp:484-490) around the unmodified M17Demodulator.h, with the
// codec2/boost consumers replaced by one hex record per frame callback.
// TEST INFRASTRUCTURE ONLY (see m17_oracle.h).
//
//   m17-demod-ref [-i] < baseband.s16 > records.txt
//   m17-demod-ref --batch FILE N_STREAMS N_SAMPLES [--jobs J] [--quiet]
//       FILE = stream-major int16 LE; one forked process per stream, J in
//       flight; prints wall seconds, Msamples/s and the record count.

#include "m17_oracle.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

extern "C" size_t m17ref_demod_once(const int16_t*, size_t, unsigned, uint32_t, m17o_record*, size_t, m17o_trace*);
extern "C" size_t m17ref_ax25_print(const uint8_t*, size_t, char*, size_t);
extern "C" int m17ref_demod_batch(const int16_t*, size_t, size_t, size_t, unsigned, m17o_record*, size_t,
                                  uint64_t*, m17o_trace*, int);

static void print_record(const m17o_record& r)
{
    static const char* names[] = {"LSF", "LICH", "STREAM", "BASIC_PACKET", "FULL_PACKET", "BERT"};
    static const int plen[] = {0, 6, 18, 26, 26, 25};
    std::printf("%u %u %llu %s cost=%d payload=", r.stream, r.seq, (unsigned long long) r.sample_pos,
                names[r.type], r.viterbi_cost);
    for (int i = 0; i != plen[r.type]; ++i) std::printf("%02x", r.payload[i]);
    std::printf(" lsf=");
    for (int i = 0; i != 30; ++i) std::printf("%02x", r.lsf[i]);
    std::printf("\n");
}

int main(int argc, char** argv)
{
    unsigned flags = 0;
    bool batch = false, quiet = false;
    int jobs = int(sysconf(_SC_NPROCESSORS_ONLN));
    std::vector<std::string> pos;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-i") flags |= M17O_INVERT;
        else if (a == "--batch") batch = true;
        else if (a == "--quiet") quiet = true;
        else if (a == "--jobs" && i + 1 < argc) jobs = std::atoi(argv[++i]);
        else if (a == "--ax25-print" && i + 1 < argc) {
            // the reference's ax25_frame + write() (ax25_frame.h:22-263) on the bytes of a file, to stdout
            std::FILE* f = std::fopen(argv[++i], "rb");
            if (!f) return 2;
            std::vector<uint8_t> bytes;
            int ch;
            while ((ch = std::fgetc(f)) != EOF) bytes.push_back(uint8_t(ch));
            std::fclose(f);
            std::vector<char> text(1 << 16);
            size_t n = m17ref_ax25_print(bytes.data(), bytes.size(), text.data(), text.size());
            std::fwrite(text.data(), 1, n < text.size() - 1 ? n : text.size() - 1, stdout);
            return 0;
        }
        else pos.push_back(a);
    }

    if (!batch) {
        std::vector<int16_t> in;
        int16_t buf[4096];
        size_t got;
        while ((got = std::fread(buf, 2, 4096, stdin)) > 0) in.insert(in.end(), buf, buf + got);
        std::vector<m17o_record> rec(in.size() / 960 + 64);
        size_t n = m17ref_demod_once(in.data(), in.size(), flags | M17O_EOF_REPEAT, 0, rec.data(), rec.size(), nullptr);
        if (n > rec.size()) n = rec.size();
        for (size_t i = 0; i != n; ++i) print_record(rec[i]);
        return 0;
    }

    if (pos.size() < 3) { std::fprintf(stderr, "usage: --batch FILE N_STREAMS N_SAMPLES\n"); return 2; }
    size_t n_streams = std::strtoull(pos[1].c_str(), nullptr, 10);
    size_t n_samples = std::strtoull(pos[2].c_str(), nullptr, 10);
    int fd = open(pos[0].c_str(), O_RDONLY);
    if (fd < 0) { std::perror("open"); return 2; }
    size_t bytes = n_streams * n_samples * 2;
    void* p = mmap(nullptr, bytes, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    if (p == MAP_FAILED) { std::perror("mmap"); return 2; }
    size_t cap = n_samples / 960 + 64;
    std::vector<m17o_record> rec(n_streams * cap);
    std::vector<uint64_t> counts(n_streams);
    auto t0 = std::chrono::steady_clock::now();
    int rc = m17ref_demod_batch(static_cast<const int16_t*>(p), n_streams, n_samples, n_samples, flags,
                                rec.data(), cap, counts.data(), nullptr, jobs);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    uint64_t total = 0;
    for (auto c : counts) total += c;
    if (!quiet)
        for (size_t s = 0; s != n_streams; ++s)
            for (size_t i = 0; i != counts[s] && i != cap; ++i) print_record(rec[s * cap + i]);
    std::fprintf(stderr, "{\"rc\": %d, \"streams\": %zu, \"samples_per_stream\": %zu, \"jobs\": %d, \"seconds\": %.6f, "
                         "\"msamples_per_s\": %.3f, \"records\": %llu}\n",
                 rc, n_streams, n_samples, jobs, dt, n_streams * n_samples / dt / 1e6, (unsigned long long) total);
    return rc ? 1 : 0;
}
