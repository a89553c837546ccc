# tools/vp9author.mk -- TEST TOOLING: the stream-authoring build of the reference decoder (tools/vp9author.c).
# The reference sources compiled by ffsubset.mk, with patched COPIES of vp9.c / vp9mvs.c / vpx_rac.h / vp89_rac.h
# (tools/vp9author_overlay.py) shadowing the originals, linked with the authoring tool.  Output: $(OUT)/ffvp9_author
# (oracle/_ref/author/, git-ignored).   make -C oracle -f ../tools/vp9author.mk
OUT := _ref/author
VP9_OVERLAY_DIR   := $(OUT)/src/libavcodec
VP9_OVERLAY_FILES := vp9 vp9mvs vp9block vp9prob vp9recon vp9lpf vp9data
VP9_OVERLAY_STAMP := $(OUT)/src/.patched
EXTRA_INC :=
include ../tools/ffsubset.mk

all: $(OUT)/ffvp9_author

$(VP9_OVERLAY_STAMP): ../tools/vp9author_overlay.py
	rm -rf $(OUT)/src && mkdir -p $(VP9_OVERLAY_DIR) $(OUT)/obj
	python3 ../tools/vp9author_overlay.py $(REF) $(OUT)/src
	for f in vp9block vp9prob vp9recon vp9lpf vp9data; do cp $(REF)/libavcodec/$$f.c $(VP9_OVERLAY_DIR)/ || exit 1; done
	cp $(REF)/libavcodec/vp9data.h $(VP9_OVERLAY_DIR)/

$(addprefix $(VP9_OVERLAY_DIR)/,$(addsuffix .c,$(VP9_OVERLAY_FILES))): $(VP9_OVERLAY_STAMP)

$(OUT)/ffvp9_author: ../tools/vp9author.c $(OUT)/libffvp9.a
	$(CC) -I$(OUT)/src $(FF_CFLAGS) $< $(OUT)/libffvp9.a -lm -o $@

.DEFAULT_GOAL := all
.PHONY: all
