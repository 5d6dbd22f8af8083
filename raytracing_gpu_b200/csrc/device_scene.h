// device_scene.h — interface between the C-ABI layer (c_api.cpp) and the CUDA side (device_*.cu).
#pragma once
#include <string>
#include <vector>
#include "scene_host.h"

struct rt_scene;

namespace rtb {

struct DeviceScene;   // defined in device_scene.cuh (CUDA translation units only)

void set_error(const std::string& e);
HostScene& host_of(rt_scene* s);
DeviceScene*& device_of(rt_scene* s);
std::vector<DeviceScene*>& replicas_of(rt_scene* s);
void device_scene_destroy(DeviceScene* d);
bool device_scene_is_current(rt_scene* s);   // a device copy exists and describes the scene as recorded now

} // namespace rtb
