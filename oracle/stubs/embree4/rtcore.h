// Minimal stand-in for <embree4/rtcore.h> so that the reference's shape classes
// (src/shapes/triangle.hpp:6-7) compile for the primitive-level cross-check.
// TEST INFRASTRUCTURE.  Embree is never called: Triangle::globalScene stays null
// (triangle.hpp:49,137), so these no-ops are unreachable at run time.
#pragma once
#include <cstddef>
#include <cstring>
#include <limits>
typedef struct RTCDeviceTy* RTCDevice;
typedef struct RTCSceneTy* RTCScene;
typedef struct RTCGeometryTy* RTCGeometry;
enum RTCGeometryType { RTC_GEOMETRY_TYPE_TRIANGLE = 0 };
enum RTCBufferType { RTC_BUFFER_TYPE_INDEX = 0, RTC_BUFFER_TYPE_VERTEX = 1 };
enum RTCFormat { RTC_FORMAT_UINT3 = 0, RTC_FORMAT_FLOAT3 = 1 };
#define RTC_INVALID_GEOMETRY_ID ((unsigned int)-1)
struct RTCRay { float org_x, org_y, org_z, tnear, dir_x, dir_y, dir_z, time, tfar; unsigned mask, id, flags; };
struct RTCHit { float Ng_x, Ng_y, Ng_z, u, v; unsigned primID, geomID, instID[1]; };
struct RTCRayHit { RTCRay ray; RTCHit hit; };
inline RTCGeometry rtcNewGeometry(RTCDevice, RTCGeometryType) { return nullptr; }
inline RTCDevice rtcGetSceneDevice(RTCScene) { return nullptr; }
inline void* rtcSetNewGeometryBuffer(RTCGeometry, RTCBufferType, unsigned, RTCFormat, size_t, size_t) { return nullptr; }
inline void rtcCommitGeometry(RTCGeometry) {}
inline unsigned rtcAttachGeometry(RTCScene, RTCGeometry) { return 0; }
inline void rtcReleaseGeometry(RTCGeometry) {}
inline void rtcCommitScene(RTCScene) {}
inline void rtcIntersect1(RTCScene, RTCRayHit*) {}
