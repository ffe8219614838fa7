// rtc_geometry_owner.h — host storage behind an rtc_geometry (include/rtc.h), whichever producer filled it
// (rtc_geometry_build on the host threads, rtc_geometry_build_device on the GPU); rtc_geometry_free deletes it.
#pragma once
#include "../../include/rtc.h"

#include <vector>

struct RtcGeometryOwner {
    std::vector<rtc_cell> cells;
    std::vector<rtc_tri_pos> tri_pos;
    std::vector<rtc_tri_data> tri_data;
};
