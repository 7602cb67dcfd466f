# model.cmake for MODEL=b200 -- the B200-native (sm_100a) BabelStream model.
#
# Same contract as every src/<model>/model.cmake of the reference (cmake/register_models.cmake):
# flags are declared with register_flag_*, setup() runs before add_executable, setup_target()
# after it.  The host class (B200Stream.cpp) is plain C++; the kernels live in
# libb200stream (CUDA, sm_100a) and are added as extra sources in setup_target(), the way the
# futhark model adds its generated C file (src/futhark/model.cmake:53-56).

register_flag_optional(CMAKE_CXX_COMPILER
        "Any CXX compiler that is supported by CMake detection, used for the host class and as nvcc's host compiler"
        "c++")

register_flag_required(CMAKE_CUDA_COMPILER
        "Path to the CUDA nvcc compiler (CUDA >= 12.8 for sm_100a)")

register_flag_optional(B200_KERNEL_DIR
        "Directory holding the kernel sources (*.cu) and include/b200_stream.h's parent"
        "${CMAKE_CURRENT_LIST_DIR}/../../csrc")

register_flag_optional(B200_INCLUDE_DIR
        "Directory holding b200_stream.h"
        "${CMAKE_CURRENT_LIST_DIR}/../../../include")

register_flag_optional(CUDA_EXTRA_FLAGS
        "Additional CUDA flags passed to nvcc"
        "")

macro(setup)
    if (POLICY CMP0104)
        cmake_policy(SET CMP0104 OLD)
    endif ()
    set(CMAKE_CXX_STANDARD 14)
    set(CMAKE_CUDA_STANDARD 17)
    enable_language(CUDA)
    # sm_100a only; -fmad=false because every FMA is written by hand (rounding parity with omp/serial)
    set(CMAKE_CUDA_FLAGS ${CMAKE_CUDA_FLAGS} "-forward-unknown-to-host-compiler"
            "-gencode arch=compute_100a,code=sm_100a" "-lineinfo" "-fmad=false" ${CUDA_EXTRA_FLAGS})
    string(REPLACE ";" " " CMAKE_CUDA_FLAGS "${CMAKE_CUDA_FLAGS}")
    wipe_gcc_style_optimisation_flags(CMAKE_CUDA_FLAGS_${BUILD_TYPE})
    include_directories(${B200_INCLUDE_DIR})
    register_link_library(dl pthread)
    message(STATUS "NVCC flags: ${CMAKE_CUDA_FLAGS} ${CMAKE_CUDA_FLAGS_${BUILD_TYPE}}")
endmacro()

macro(setup_target NAME)
    file(GLOB B200_KERNEL_SOURCES "${B200_KERNEL_DIR}/*.cu")
    target_sources(${NAME} PUBLIC ${B200_KERNEL_SOURCES})
    set_source_files_properties(${B200_KERNEL_SOURCES} PROPERTIES COMPILE_OPTIONS "-O3")
endmacro()
