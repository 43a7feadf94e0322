// oracle/refprobe_stubs.cpp — TEST INFRASTRUCTURE ONLY.
// Link-time stand-ins for the handful of reference symbols the hot-path templates reference but never
// reach from the probe's entry points (logging, stats registry, a spectrum type the probe does not
// construct).  Everything on the measured/parity path is the reference's own object code.
#include <pbrt/options.h>
#include <pbrt/samplers.h>
#include <pbrt/util/bluenoise.h>
#include <pbrt/util/log.h>
#include <pbrt/util/parallel.h>
#include <pbrt/util/progressreporter.h>
#include <pbrt/util/error.h>
#include <pbrt/util/print.h>
#include <pbrt/util/spectrum.h>
#include <pbrt/util/stats.h>

#include <cstdio>
#include <cstdlib>

namespace pbrt {
void LogFatal(LogLevel, const char *file, int line, const char *s) {
    std::fprintf(stderr, "refprobe: reference CHECK failed at %s:%d: %s\n", file, line, s);
    std::abort();
}
// NanoVDBMedium's constructor (media.cpp:511-616) logs at Verbose level and fills the majorant grid
// with ParallelFor; the probe runs that loop body serially (scheduling only, same per-cell code).
namespace logging { LogLevel logLevel = LogLevel::Error; }
void Log(LogLevel, const char *file, int line, const char *s) { std::fprintf(stderr, "refprobe: %s:%d: %s\n", file, line, s); }
void ParallelFor(int64_t start, int64_t end, std::function<void(int64_t, int64_t)> func) { func(start, end); }
// src/graph prints progress bars and exits on errors; neither is arithmetic of the path.
ProgressReporter::ProgressReporter(int64_t totalWork, std::string title, bool, bool)
    : totalWork(totalWork), title(std::move(title)), quiet(true), workDone(0), exitThread(true), printPlusStatus(0) {}
ProgressReporter::~ProgressReporter() {}
void ProgressReporter::Done() {}
void ErrorExit(const FileLoc *, const char *message) {
    std::fprintf(stderr, "refprobe: reference ErrorExit: %s\n", message);
    std::abort();
}
// ZSobolSampler's constructor warns about sample counts that are not powers of two (samplers.h:231-234)
void Warning(const FileLoc *, const char *) {}
int RunningThreads() { return 1; }
StatRegisterer::StatRegisterer(AccumFunc, PixelAccumFunc) {}
void StatsAccumulator::ReportMemoryCounter(const char *, int64_t) {}
void StatsAccumulator::ReportPercentage(const char *, int64_t, int64_t) {}
PBRTOptions *Options = nullptr;  // options.cpp:14 (the probe points it at a local PBRTOptions)
// bluenoise.cpp is one of the blobs missing from the reference mount (.MISSING_LARGE_BLOBS); only the
// PMJ02BN sampler reads it, which the probe never instantiates.
const uint16_t BlueNoiseTextures[NumBlueNoiseTextures][BlueNoiseResolution][BlueNoiseResolution] = {};
Float PiecewiseLinearSpectrum::operator()(Float) const {
    std::fprintf(stderr, "refprobe: PiecewiseLinearSpectrum is not part of the probe\n");
    std::abort();
}
namespace detail {
// only reached from ToString()/StringPrintf on error paths
std::string FloatToString(float v) { return std::to_string(v); }
std::string DoubleToString(double v) { return std::to_string(v); }
void stringPrintfRecursive(std::string *s, const char *fmt) { *s += fmt; }
std::string copyToFormatString(const char **fmt_ptr, std::string *s) {
    *s += *fmt_ptr;
    while (**fmt_ptr) ++*fmt_ptr;
    return "%s";
}
}  // namespace detail
}  // namespace pbrt
