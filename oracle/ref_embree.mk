# Builds the reference's vendored Embree 3.6.1 (dependencies/embree @ d17aec1) as ONE shared
# object, straight from the sources where they lie under /root/reference, with g++ only
# (no cmake, no copy of any reference source into this repo).  Outputs go to oracle/_ref/.
#
# TEST INFRASTRUCTURE ONLY: the product (libwpt.so) never links or loads this.
#
# Single-ISA build: every file is compiled with Embree's own AVX2 flag set
# (common/cmake/gnu.cmake FLAGS_AVX2), so `isa` == avx2 and BVH8.Triangle4 + Moeller-Trumbore
# is the default accel (kernels/common/scene.cpp:120-160), exactly the configuration SURVEY §8c
# pins.  Geometry types other than triangles, ray packets and ISPC are compiled out, mirroring the
# reference's own (commented) option block CMakeLists.txt:13-29.
#
# The three headers Embree's CMake would generate (kernels/config.h, kernels/hash.h,
# include/embree3/rtcore_config.h) are produced here with sed from the *.in templates into
# oracle/_ref/gen/ and found through -I.

REF      ?= /root/reference/dependencies/embree
OUT      ?= $(CURDIR)/_ref
OBJ      := $(OUT)/obj
GEN      := $(OUT)/gen
CXX      := /usr/bin/g++
ISAFLAGS := -mf16c -mavx2 -mfma -mlzcnt -mbmi -mbmi2
CXXFLAGS := -std=c++11 -O3 -DNDEBUG -fPIC -fvisibility=hidden -fvisibility-inlines-hidden \
            -fno-strict-aliasing -fno-tree-vectorize -w $(ISAFLAGS) \
            -DTASKING_INTERNAL -DEMBREE_LOWEST_ISA -DRTC_EXPORT_API -D__TARGET_AVX2__ \
            -I$(GEN) -I$(GEN)/sub -I$(REF)/include -I$(REF)

SRCS := \
  common/sys/sysinfo.cpp common/sys/alloc.cpp common/sys/filename.cpp common/sys/library.cpp \
  common/sys/thread.cpp common/sys/string.cpp common/sys/regression.cpp common/sys/mutex.cpp \
  common/sys/condition.cpp common/sys/barrier.cpp \
  common/math/constants.cpp common/simd/sse.cpp \
  common/lexers/stringstream.cpp common/lexers/tokenstream.cpp \
  common/tasking/taskschedulerinternal.cpp \
  common/algorithms/parallel_for.cpp common/algorithms/parallel_reduce.cpp \
  common/algorithms/parallel_prefix_sum.cpp common/algorithms/parallel_for_for.cpp \
  common/algorithms/parallel_for_for_prefix_sum.cpp common/algorithms/parallel_partition.cpp \
  common/algorithms/parallel_sort.cpp common/algorithms/parallel_set.cpp \
  common/algorithms/parallel_map.cpp common/algorithms/parallel_filter.cpp \
  kernels/common/device.cpp kernels/common/stat.cpp kernels/common/acceln.cpp \
  kernels/common/accelset.cpp kernels/common/state.cpp kernels/common/rtcore.cpp \
  kernels/common/rtcore_builder.cpp kernels/common/scene.cpp kernels/common/alloc.cpp \
  kernels/common/geometry.cpp kernels/common/scene_user_geometry.cpp \
  kernels/common/scene_instance.cpp kernels/common/scene_triangle_mesh.cpp \
  kernels/common/scene_quad_mesh.cpp kernels/common/scene_curves.cpp \
  kernels/common/scene_line_segments.cpp kernels/common/scene_grid_mesh.cpp \
  kernels/common/scene_points.cpp \
  kernels/subdiv/bezier_curve.cpp kernels/subdiv/bspline_curve.cpp kernels/subdiv/catmullrom_curve.cpp \
  kernels/geometry/primitive4.cpp kernels/geometry/primitive8.cpp \
  kernels/geometry/instance_intersector.cpp kernels/geometry/curve_intersector_virtual.cpp \
  kernels/builders/primrefgen.cpp \
  kernels/bvh/bvh.cpp kernels/bvh/bvh_statistics.cpp kernels/bvh/bvh4_factory.cpp kernels/bvh/bvh8_factory.cpp \
  kernels/bvh/bvh_rotate.cpp kernels/bvh/bvh_refit.cpp kernels/bvh/bvh_builder.cpp \
  kernels/bvh/bvh_builder_hair.cpp kernels/bvh/bvh_builder_hair_mb.cpp kernels/bvh/bvh_builder_morton.cpp \
  kernels/bvh/bvh_builder_sah.cpp kernels/bvh/bvh_builder_sah_spatial.cpp kernels/bvh/bvh_builder_sah_mb.cpp \
  kernels/bvh/bvh_builder_twolevel.cpp \
  kernels/bvh/bvh_intersector1_bvh4.cpp kernels/bvh/bvh_intersector1_bvh8.cpp

OBJS := $(patsubst %.cpp,$(OBJ)/%.o,$(SRCS))

all: $(OUT)/libembree_ref.so

$(GEN)/stamp:
	mkdir -p $(GEN)/sub
	sed -e 's/#cmakedefine EMBREE_GEOMETRY_TRIANGLE/#define EMBREE_GEOMETRY_TRIANGLE/' \
	    -e 's/#cmakedefine EMBREE_FILTER_FUNCTION/#define EMBREE_FILTER_FUNCTION/' \
	    -e 's/#cmakedefine EMBREE_GEOMETRY_USER/#define EMBREE_GEOMETRY_USER/' \
	    -e 's/#cmakedefine \(.*\)/\/* #undef \1 *\//' \
	    -e 's/@EMBREE_CURVE_SELF_INTERSECTION_AVOIDANCE_FACTOR@/2.0/' \
	    $(REF)/kernels/config.h.in > $(GEN)/config.h
	sed -e 's/@EMBREE_HASH@/d17aec1/' $(REF)/kernels/hash.h.in > $(GEN)/hash.h
	sed -e 's/@EMBREE_VERSION_MAJOR@/3/g' -e 's/@EMBREE_VERSION_MINOR@/6/g' \
	    -e 's/@EMBREE_VERSION_PATCH@/1/g' -e 's/@EMBREE_VERSION_NUMBER@/30601/g' \
	    -e 's/@EMBREE_VERSION_NOTE@//g' -e 's/@EMBREE_MAX_INSTANCE_LEVEL_COUNT@/1/g' \
	    -e 's/#cmakedefine EMBREE_STATIC_LIB/#define EMBREE_STATIC_LIB/' \
	    -e 's/#cmakedefine \(.*\)/\/* #undef \1 *\//' \
	    $(REF)/kernels/rtcore_config.h.in > $(GEN)/rtcore_config.h
	touch $@

$(OBJ)/%.o: $(REF)/%.cpp $(GEN)/stamp
	@mkdir -p $(dir $@)
	@echo "  CXX $<"; $(CXX) $(CXXFLAGS) -c $< -o $@

# embree_ref_wrap.cpp is OUR thin C wrapper (scene create / rtcIntersect1 loop) over the public API
$(OBJ)/embree_ref_wrap.o: embree_ref_wrap.cpp $(GEN)/stamp
	@mkdir -p $(dir $@)
	$(CXX) -std=c++17 -O3 -fPIC -fopenmp $(ISAFLAGS) -I$(GEN) -I$(REF)/include -I. -c $< -o $@

$(OUT)/libembree_ref.so: $(OBJS) $(OBJ)/embree_ref_wrap.o
	@echo "  LD $@"; $(CXX) -shared -fopenmp -o $@ $^ -lpthread -ldl

clean:
	rm -rf $(OUT)
.PHONY: all clean
