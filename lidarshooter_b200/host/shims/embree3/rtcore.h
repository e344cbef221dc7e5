// shim: the one Embree 3 type in ITracer's signature (SRC/ITracer.hpp:50); enumerator values of embree3/rtcore_geometry.h
#pragma once
enum RTCGeometryType {
    RTC_GEOMETRY_TYPE_TRIANGLE = 0,
    RTC_GEOMETRY_TYPE_QUAD = 1,
    RTC_GEOMETRY_TYPE_GRID = 2,
    RTC_GEOMETRY_TYPE_SUBDIVISION = 8
};
