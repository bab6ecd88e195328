# r2iq_cuda.cmake -- build switch that puts the B200 operator behind the reference's own class names.
#
# In the reference tree, Core/CMakeLists.txt gains three lines (nothing else changes anywhere):
#
#     include(${R2IQ_CUDA_ROOT}/extio_sddc_b200/host/r2iq_cuda.cmake)     # after the SRC list is final (before line 81)
#     r2iq_cuda_sources(SRC)                                               # swaps the operator's sources
#     add_library(SDDC_CORE STATIC ${SRC})                                 # the existing line 81
#     r2iq_cuda_link(SDDC_CORE)                                            # after add_library
#
# and the tree is configured with  -DSDDC_R2IQ_CUDA=ON -DR2IQ_CUDA_ROOT=/path/to/this/repo
# (optionally -DSDDC_FFT_BACKEND_CUDA=ON to make CudaFFTBackend the compiled-in getFFTBackend()).
#
# With the switch ON, Core/fft_mt_r2iq.cpp and its per-ISA workers (Core/CMakeLists.txt:44-72) leave the build and
# host/fft_mt_r2iq_cuda.cpp + host/cuda_r2iq.cpp enter it: `new fft_mt_r2iq()` in Core/RadioHandler.cpp:94-95 -- the
# operator every application ends up with (ExtIO_sddc, libsddc, SoapySDDC pass a null operator) -- is then the GPU
# one.  With the switch OFF the build is the reference's, byte for byte.
option(SDDC_R2IQ_CUDA "Run the fft_mt_r2iq down-converter on a B200 through libr2iq_cuda.so" OFF)
option(SDDC_FFT_BACKEND_CUDA "Make CudaFFTBackend the compiled-in FFT backend (Core/fft_backend.h)" OFF)
set(R2IQ_CUDA_ROOT "${CMAKE_CURRENT_LIST_DIR}/../.." CACHE PATH "Root of the extio_sddc_b200 repository")

macro(r2iq_cuda_sources SRC_VAR)
    if(SDDC_R2IQ_CUDA)
        list(FILTER ${SRC_VAR} EXCLUDE REGEX "fft_mt_r2iq(_def|_avx|_avx2|_avx512|_neon)?\\.cpp$")
        list(APPEND ${SRC_VAR}
             "${R2IQ_CUDA_ROOT}/extio_sddc_b200/host/fft_mt_r2iq_cuda.cpp"
             "${R2IQ_CUDA_ROOT}/extio_sddc_b200/host/cuda_r2iq.cpp")
        message(STATUS "r2iq operator: CUDA (libr2iq_cuda.so from ${R2IQ_CUDA_ROOT})")
    endif()
    if(SDDC_FFT_BACKEND_CUDA)
        list(FILTER ${SRC_VAR} EXCLUDE REGEX "fft_backend_(fftw|mkl|accelerate)\\.cpp$")
        list(APPEND ${SRC_VAR} "${R2IQ_CUDA_ROOT}/extio_sddc_b200/host/cuda_fft_backend.cpp")
        message(STATUS "FFT backend: CUDA")
    endif()
endmacro()

function(r2iq_cuda_link TARGET)
    if(NOT SDDC_R2IQ_CUDA AND NOT SDDC_FFT_BACKEND_CUDA)
        return()
    endif()
    set(_lib_dir "${R2IQ_CUDA_ROOT}/extio_sddc_b200/lib")
    if(NOT EXISTS "${_lib_dir}/libr2iq_cuda.so")
        message(FATAL_ERROR "libr2iq_cuda.so not found in ${_lib_dir}: run `python -m extio_sddc_b200.build` there first "
                            "(the operator is CUDA-only, there is no CPU fallback to link instead)")
    endif()
    target_compile_definitions(${TARGET} PUBLIC R2IQ_USE_REFERENCE_HEADERS)
    if(SDDC_FFT_BACKEND_CUDA)
        target_compile_definitions(${TARGET} PRIVATE R2IQ_FFT_BACKEND_FACTORY)
    endif()
    target_include_directories(${TARGET} PUBLIC "${R2IQ_CUDA_ROOT}/include" "${R2IQ_CUDA_ROOT}/extio_sddc_b200/host")
    target_link_directories(${TARGET} PUBLIC "${_lib_dir}")
    target_link_libraries(${TARGET} PUBLIC r2iq_cuda)
    set_property(TARGET ${TARGET} APPEND PROPERTY INTERFACE_LINK_OPTIONS "-Wl,-rpath,${_lib_dir}")
endfunction()
