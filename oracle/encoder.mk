# TEST INFRASTRUCTURE ONLY — whole-encoder drop-in check (tests/test_encoder_dropin.py).
#
#   make -f oracle/encoder.mk        (only where $(REF) = /root/reference exists)
#
# Compiles the reference ENCODER (library + command-line application) from its own sources where they lie, C only (no
# NASM / SIMD objects: the RTCD tables then hold the C kernels, aom_dsp_rtcd.c:1199-1214), with plain gcc and this
# file instead of the reference's cmake, and links it twice:
#   oracle/_ref/enc/SvtAv1EncApp_ref   the unmodified encoder
#   oracle/_ref/enc/SvtAv1EncApp_gpu   the same objects + svt-av1-psy_b200/shim/svtme_wrap.c, linked with
#                                      -Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0
#                                      against libsvtme_b200.so: every open-loop ME, temporal-filter ME and dynamic-GOP
#                                      detector search of the encode runs on the GPU
#   oracle/_ref/enc/SvtAv1EncApp_gpu_batched   the same again, with ONE object replaced: pd_process.c patched by
#                                      oracle/pd_process_submit.patch (applied to a COPY under oracle/_ref/enc/, the reference
#                                      tree is not touched): Picture Decision hands every mini-GOP to
#                                      svtme_shim_submit_pictures before posting its ME tasks = the throughput path
#                                      (one GPU batch and one result download per mini-GOP)
# All three must write the same bitstream. The one header cmake would generate (EbVersion.h: two string defines from
# EbVersion.h.in) is written here.
REF   ?= /root/reference
RS    := $(REF)/Source
HERE  := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))
PKG   := $(HERE)../svt-av1-psy_b200
OUT   := $(HERE)_ref/enc
CC    ?= gcc
CFLAGS := -O2 -g0 -w -fPIC -DEXCLUDE_HASH=1 -DREPRODUCIBLE_BUILDS=1 -DEN_AVX512_SUPPORT=0 -D_GNU_SOURCE
INC   := -I$(OUT) -I$(RS)/API -I$(RS)/Lib/Codec -I$(RS)/Lib/C_DEFAULT -I$(RS)/Lib/Globals -I$(REF)/third_party/fastfeat \
         -I$(REF)/third_party/aom_dsp/inc -I$(REF)/third_party/safestringlib -I$(REF)

LIB_SRCS := $(wildcard $(RS)/Lib/Codec/*.c) $(wildcard $(RS)/Lib/C_DEFAULT/*.c) $(wildcard $(RS)/Lib/Globals/*.c) \
            $(wildcard $(REF)/third_party/fastfeat/*.c)
APP_SRCS := $(wildcard $(RS)/App/*.c) $(wildcard $(REF)/third_party/safestringlib/*.c)
LIB_OBJS := $(patsubst $(REF)/%.c,$(OUT)/obj/%.o,$(LIB_SRCS))
APP_OBJS := $(patsubst $(REF)/%.c,$(OUT)/obj/%.o,$(APP_SRCS))

.PHONY: all clean
all: $(OUT)/SvtAv1EncApp_ref $(OUT)/SvtAv1EncApp_gpu $(OUT)/SvtAv1EncApp_gpu_batched

$(OUT)/EbVersion.h:
	@mkdir -p $(OUT)
	@printf '#define SVT_AV1_CVS_VERSION "2.3.0-A (oracle/encoder.mk)"\n#define SVT_AV1_PSY_RELEASE "A"\n' > $@

$(OUT)/obj/%.o: $(REF)/%.c | $(OUT)/EbVersion.h
	@mkdir -p $(dir $@)
	@$(CC) $(CFLAGS) $(INC) -c $< -o $@

$(OUT)/svtme_wrap.o: $(PKG)/shim/svtme_wrap.c $(PKG)/shim/svtme_wrap.h $(HERE)../include/svtme_b200.h | $(OUT)/EbVersion.h
	@$(CC) $(CFLAGS) -Wall $(INC) -I$(HERE)../include -I$(PKG)/shim -c $< -o $@

$(OUT)/SvtAv1EncApp_ref: $(LIB_OBJS) $(APP_OBJS)
	@$(CC) -o $@ $(APP_OBJS) $(LIB_OBJS) -lm -lpthread

$(OUT)/SvtAv1EncApp_gpu: $(LIB_OBJS) $(APP_OBJS) $(OUT)/svtme_wrap.o $(PKG)/libsvtme_b200.so
	@$(CC) -o $@ $(APP_OBJS) $(LIB_OBJS) $(OUT)/svtme_wrap.o \
	    -Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0 \
	    -L$(PKG) -lsvtme_b200 -Wl,-rpath,'$$ORIGIN/../../../svt-av1-psy_b200' -lm -lpthread

PD_OBJ := $(OUT)/obj/Source/Lib/Codec/pd_process.o
$(OUT)/pd_process_submit.c: $(RS)/Lib/Codec/pd_process.c $(HERE)pd_process_submit.patch | $(OUT)/EbVersion.h
	@patch -s -o $@ $< $(HERE)pd_process_submit.patch

$(OUT)/pd_process_submit.o: $(OUT)/pd_process_submit.c
	@$(CC) $(CFLAGS) $(INC) -c $< -o $@

$(OUT)/SvtAv1EncApp_gpu_batched: $(LIB_OBJS) $(APP_OBJS) $(OUT)/svtme_wrap.o $(OUT)/pd_process_submit.o $(PKG)/libsvtme_b200.so
	@$(CC) -o $@ $(APP_OBJS) $(filter-out $(PD_OBJ),$(LIB_OBJS)) $(OUT)/pd_process_submit.o $(OUT)/svtme_wrap.o \
	    -Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0 \
	    -L$(PKG) -lsvtme_b200 -Wl,-rpath,'$$ORIGIN/../../../svt-av1-psy_b200' -lm -lpthread

clean:
	rm -rf $(OUT)
