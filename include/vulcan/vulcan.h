// vulcan/vulcan.h - umbrella include of the fusion hot path (the reference generates its
// vulcan.h from vulcan.h.in with every header of the library, CMakeLists.txt:136-143).
#pragma once

#include <vulcan/block.h>
#include <vulcan/buffer.h>
#include <vulcan/color_integrator.h>
#include <vulcan/depth_integrator.h>
#include <vulcan/depth_tracker.h>
#include <vulcan/device.h>
#include <vulcan/exception.h>
#include <vulcan/exporter.h>
#include <vulcan/extractor.h>
#include <vulcan/frame.h>
#include <vulcan/frame_pipeline.h>
#include <vulcan/hash.h>
#include <vulcan/image.h>
#include <vulcan/integrator.h>
#include <vulcan/light.h>
#include <vulcan/light_integrator.h>
#include <vulcan/math.h>
#include <vulcan/matrix.h>
#include <vulcan/mesh.h>
#include <vulcan/projection.h>
#include <vulcan/pyramid_tracker.h>
#include <vulcan/sharded_tracer.h>
#include <vulcan/sharded_pipeline.h>
#include <vulcan/tracer.h>
#include <vulcan/tracer.cuh>
#include <vulcan/tracker.h>
#include <vulcan/transform.h>
#include <vulcan/types.h>
#include <vulcan/volume.h>
#include <vulcan/voxel.h>
