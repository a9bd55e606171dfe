// oracle/qt_shim — TEST INFRASTRUCTURE ONLY.
// Minimal stand-ins for the few Qt names the reference's pebblelib sources touch, so that the
// UNMODIFIED reference .cpp files under /root/reference compile without Qt. No arithmetic here.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <algorithm>
typedef int8_t qint8;   typedef uint8_t quint8;
typedef int16_t qint16; typedef uint16_t quint16;
typedef int32_t qint32; typedef uint32_t quint32;
typedef int64_t qint64; typedef uint64_t quint64;
typedef double qreal;
#define Q_DECL_EXPORT __attribute__((visibility("default")))
#define Q_DECL_IMPORT
#define Q_UNUSED(x) (void)x;
#define Q_ASSERT(x) ((void)0)
#define Q_OBJECT
#define signals public
#define slots
#define emit
template <typename T> inline const T &qBound(const T &lo, const T &v, const T &hi) {
    return std::max(lo, std::min(hi, v));
}
template <typename T> inline const T &qMin(const T &a, const T &b) { return std::min(a, b); }
template <typename T> inline const T &qMax(const T &a, const T &b) { return std::max(a, b); }
