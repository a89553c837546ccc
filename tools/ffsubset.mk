# tools/ffsubset.mk -- compile the slice of the reference tree that the "vp9" decoder needs,
# straight from $(REF) with plain gcc; never runs the reference's configure/ffbuild.
# Includers set OUT (output dir) and may set VP9_OVERLAY_DIR: a directory holding patched copies
# of libavcodec files that shadow the reference ones (used by the vp9cuda host build only).
#
# -fwrapv: the synthetic streams contain out-of-range coefficients on purpose; signed overflow in
# the reference's integer butterflies is then two's-complement wrap instead of undefined.
REF    ?= /root/reference
TOOLS  := $(dir $(lastword $(MAKEFILE_LIST)))
INC    := $(OUT)/inc
CC     ?= gcc
FF_CFLAGS := -std=c17 -O2 -fPIC -fwrapv -w -pthread -fno-math-errno -fno-signed-zeros \
          -D_ISOC11_SOURCE -D_FILE_OFFSET_BITS=64 -D_LARGEFILE_SOURCE -D_POSIX_C_SOURCE=200112 \
          -D_XOPEN_SOURCE=600 -D_DEFAULT_SOURCE -DHAVE_AV_CONFIG_H \
          -I$(INC) -I$(REF) -I$(REF)/libavcodec $(EXTRA_INC)

# the hot path and the decoder around it (reference libavcodec/Makefile:811-813)
AVC_VP9 := vp9 vp9block vp9data vp9dsp vp9dsp_8bpp vp9dsp_10bpp vp9dsp_12bpp vp9lpf vp9mvs \
           vp9prob vp9recon vpx_rac videodsp
# generic libavcodec the decoder is driven through
AVC_GEN := allcodecs avcodec decode utils options packet bsf bitstream_filters \
           bsf/vp9_superframe_split bsf/null cbs cbs_vp9 codec_desc codec_par get_buffer \
           pthread pthread_frame pthread_slice threadprogress profiles mathtables parsers parser \
           version encode raw imgconvert exif tiff_common lcevcdec
AVU     := mem log frame buffer dict opt eval parseutils samplefmt channel_layout pixdesc imgutils \
           bprint avstring error rational mathematics time random_seed cpu hwcontext side_data md5 \
           crc intmath log2_tab reverse slicethread refstruct film_grain_params \
           mastering_display_metadata stereo3d utils sha fifo display spherical \
           hdr_dynamic_metadata hdr_dynamic_vivid_metadata dovi_meta downmix_info \
           ambient_viewing_environment detection_bbox video_enc_params timecode timecode_internal \
           tdrdi iamf container_fifo executor timestamp file_open avsscanf base64

OBJ_AVC := $(foreach f,$(AVC_VP9) $(AVC_GEN),$(OUT)/obj/avc_$(subst /,_,$(f)).o)
OBJ_AVU := $(foreach f,$(AVU),$(OUT)/obj/avu_$(f).o)

$(INC)/config.h: $(TOOLS)gen_ffconfig.py
	mkdir -p $(INC) $(OUT)/obj
	python3 $(TOOLS)gen_ffconfig.py $(REF) $(INC)

# a file present in VP9_OVERLAY_DIR shadows the reference file of the same name
ffsrc = $(if $(and $(VP9_OVERLAY_DIR),$(filter $(1),$(VP9_OVERLAY_FILES))),$(VP9_OVERLAY_DIR)/$(1).c,$(REF)/libavcodec/$(1).c)

define avc_rule
$(OUT)/obj/avc_$(subst /,_,$(1)).o: $(call ffsrc,$(1)) $(INC)/config.h $(VP9_OVERLAY_STAMP)
	$(CC) $(FF_CFLAGS) -DBUILDING_avcodec -c $$< -o $$@
endef
$(foreach f,$(AVC_VP9) $(AVC_GEN),$(eval $(call avc_rule,$(f))))

$(OUT)/obj/avu_%.o: $(REF)/libavutil/%.c $(INC)/config.h
	$(CC) $(FF_CFLAGS) -DBUILDING_avutil -c $< -o $@

$(OUT)/libffvp9.a: $(OBJ_AVC) $(OBJ_AVU) $(EXTRA_OBJS)
	rm -f $@ && ar rcs $@ $^
