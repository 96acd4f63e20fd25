// Host-side formatter for master_car_statistics.csv rows (plain C++,
// no device work).  Byte-compatible with append_to_master_csv (cvs_erosion.py:232-265),
// i.e. with what pandas writes for
//   frame, car_id, matched_bbox_id, total_points, points_inside_bbox, points_outside_bbox,
//   round(inside_percentage, 2), round(outside_percentage, 2), is_matched, timestamp
// Python's round(x, 2) is the double nearest to the correctly rounded two-decimal string
// of x, and repr() of that double is that string with trailing zeros removed down to one
// decimal.  glibc's "%.2f" performs the same exact, round-half-even decimal conversion.
#include <stdio.h>
#include <string.h>
#include "lidarfuse.h"

static int fmt_pct(double x, char* buf) {
    int n = snprintf(buf, 32, "%.2f", x);
    while (n > 0 && buf[n - 1] == '0' && buf[n - 2] != '.') buf[--n] = 0;
    return n;
}

extern "C" int64_t lf_format_csv_rows(const int64_t* frame, const int32_t* car_id, const int32_t* matched_bbox_id,
                                      const int64_t* total, const int64_t* inside, int64_t n, const char* timestamp,
                                      int with_header, char* out, int64_t cap) {
    if (n < 0 || !timestamp || (n > 0 && (!frame || !car_id || !matched_bbox_id || !total || !inside))) return LF_E_ARG;
    static const char* HDR = "frame,car_id,matched_bbox_id,total_points,points_inside_bbox,points_outside_bbox,"
                             "inside_percentage,outside_percentage,is_matched,timestamp\n";
    int64_t pos = 0;
    char line[512], a[32], b[32];
    if (with_header) {
        const int64_t l = (int64_t)strlen(HDR);
        if (out && pos + l <= cap) memcpy(out + pos, HDR, (size_t)l);
        pos += l;
    }
    for (int64_t i = 0; i < n; ++i) {
        const long long t = total[i], in = inside[i], outp = t - in;
        // inside_percentage = (inside / total) * 100   (cvs_erosion.py:205-206), Python floats
        const double pi = t ? ((double)in / (double)t) * 100.0 : 0.0;
        const double po = t ? ((double)outp / (double)t) * 100.0 : 0.0;
        fmt_pct(pi, a);
        fmt_pct(po, b);
        const int l = snprintf(line, sizeof(line), "%lld,%d,%d,%lld,%lld,%lld,%s,%s,%s,%s\n", (long long)frame[i],
                               car_id[i], matched_bbox_id[i], t, in, outp, a, b,
                               matched_bbox_id[i] >= 0 ? "True" : "False", timestamp);
        if (l < 0 || l >= (int)sizeof(line)) return LF_E_ARG;
        if (out && pos + l <= cap) memcpy(out + pos, line, (size_t)l);
        pos += l;
    }
    return pos;
}
