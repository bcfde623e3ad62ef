/*
 * jpeglib.h -- minimal re-declaration of the PUBLIC libjpeg v6.2 ABI ("libjpeg62", JPEG_LIB_VERSION 62)
 * as exported by libjpeg-turbo 3.1.x binaries.  TEST INFRASTRUCTURE ONLY.
 *
 * Neither the reference tree nor this image carries libjpeg-turbo's headers or sources (SURVEY.md
 * fact 2), only binaries (Pillow bundles 3.1.4.1).  This file declares just what the reference's glue
 * (extensions/libjpeg_turbo/jpeg_mem.cpp, jpeg_handle.cpp) touches, laid out per the published ABI,
 * so that glue can be compiled from where it lies and linked against that binary by
 * oracle/build_ref.sh.  jpeg_CreateDecompress() checks both the version and
 * sizeof(struct jpeg_decompress_struct) at run time, which catches layout mistakes.
 */
#ifndef B200_TURBO_ABI_JPEGLIB_H
#define B200_TURBO_ABI_JPEGLIB_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define JPEG_LIB_VERSION 62
#define LIBJPEG_TURBO_VERSION 3.1.x
#define DCTSIZE 8
#define DCTSIZE2 64
#define NUM_QUANT_TBLS 4
#define NUM_HUFF_TBLS 4
#define NUM_ARITH_TBLS 16
#define MAX_COMPS_IN_SCAN 4
#define MAX_SAMP_FACTOR 4
#define C_MAX_BLOCKS_IN_MCU 10
#define D_MAX_BLOCKS_IN_MCU 10

typedef unsigned char JSAMPLE;
typedef short JCOEF;
typedef unsigned char JOCTET;
typedef unsigned char UINT8;
typedef unsigned short UINT16;
typedef unsigned int JDIMENSION;
typedef int boolean;
#ifndef FALSE
#define FALSE 0
#endif
#ifndef TRUE
#define TRUE 1
#endif

typedef JSAMPLE* JSAMPROW;
typedef JSAMPROW* JSAMPARRAY;
typedef JSAMPARRAY* JSAMPIMAGE;
typedef JCOEF JBLOCK[DCTSIZE2];
typedef JBLOCK* JBLOCKROW;
typedef JBLOCKROW* JBLOCKARRAY;

typedef struct JQUANT_TBL JQUANT_TBL;
typedef struct JHUFF_TBL JHUFF_TBL;
typedef struct jpeg_component_info jpeg_component_info;
typedef struct jpeg_marker_struct* jpeg_saved_marker_ptr;

typedef enum {
    JCS_UNKNOWN, JCS_GRAYSCALE, JCS_RGB, JCS_YCbCr, JCS_CMYK, JCS_YCCK, JCS_EXT_RGB, JCS_EXT_RGBX, JCS_EXT_BGR, JCS_EXT_BGRX,
    JCS_EXT_XBGR, JCS_EXT_XRGB, JCS_EXT_RGBA, JCS_EXT_BGRA, JCS_EXT_ABGR, JCS_EXT_ARGB, JCS_RGB565
} J_COLOR_SPACE;
typedef enum { JDCT_ISLOW, JDCT_IFAST, JDCT_FLOAT } J_DCT_METHOD;
#define JDCT_DEFAULT JDCT_ISLOW
#define JDCT_FASTEST JDCT_IFAST
typedef enum { JDITHER_NONE, JDITHER_ORDERED, JDITHER_FS } J_DITHER_MODE;

typedef struct jpeg_common_struct* j_common_ptr;
typedef struct jpeg_compress_struct* j_compress_ptr;
typedef struct jpeg_decompress_struct* j_decompress_ptr;

#define jpeg_common_fields \
    struct jpeg_error_mgr* err; \
    struct jpeg_memory_mgr* mem; \
    struct jpeg_progress_mgr* progress; \
    void* client_data; \
    boolean is_decompressor; \
    int global_state

struct jpeg_common_struct {
    jpeg_common_fields;
};

/* Only the head of the compression object is needed (the glue's memory destination manager is
 * compiled but never driven here). */
struct jpeg_compress_struct {
    jpeg_common_fields;
    struct jpeg_destination_mgr* dest;
};

struct jpeg_decompress_struct {
    jpeg_common_fields;
    struct jpeg_source_mgr* src;
    JDIMENSION image_width;
    JDIMENSION image_height;
    int num_components;
    J_COLOR_SPACE jpeg_color_space;
    J_COLOR_SPACE out_color_space;
    unsigned int scale_num, scale_denom;
    double output_gamma;
    boolean buffered_image;
    boolean raw_data_out;
    J_DCT_METHOD dct_method;
    boolean do_fancy_upsampling;
    boolean do_block_smoothing;
    boolean quantize_colors;
    J_DITHER_MODE dither_mode;
    boolean two_pass_quantize;
    int desired_number_of_colors;
    boolean enable_1pass_quant;
    boolean enable_external_quant;
    boolean enable_2pass_quant;
    JDIMENSION output_width;
    JDIMENSION output_height;
    int out_color_components;
    int output_components;
    int rec_outbuf_height;
    int actual_number_of_colors;
    JSAMPARRAY colormap;
    JDIMENSION output_scanline;
    int input_scan_number;
    JDIMENSION input_iMCU_row;
    int output_scan_number;
    JDIMENSION output_iMCU_row;
    int (*coef_bits)[DCTSIZE2];
    JQUANT_TBL* quant_tbl_ptrs[NUM_QUANT_TBLS];
    JHUFF_TBL* dc_huff_tbl_ptrs[NUM_HUFF_TBLS];
    JHUFF_TBL* ac_huff_tbl_ptrs[NUM_HUFF_TBLS];
    int data_precision;
    jpeg_component_info* comp_info;
    boolean progressive_mode;
    boolean arith_code;
    UINT8 arith_dc_L[NUM_ARITH_TBLS];
    UINT8 arith_dc_U[NUM_ARITH_TBLS];
    UINT8 arith_ac_K[NUM_ARITH_TBLS];
    unsigned int restart_interval;
    boolean saw_JFIF_marker;
    UINT8 JFIF_major_version;
    UINT8 JFIF_minor_version;
    UINT8 density_unit;
    UINT16 X_density;
    UINT16 Y_density;
    boolean saw_Adobe_marker;
    UINT8 Adobe_transform;
    boolean CCIR601_sampling;
    jpeg_saved_marker_ptr marker_list;
    int max_h_samp_factor;
    int max_v_samp_factor;
    int min_DCT_scaled_size;
    JDIMENSION total_iMCU_rows;
    JSAMPLE* sample_range_limit;
    int comps_in_scan;
    jpeg_component_info* cur_comp_info[MAX_COMPS_IN_SCAN];
    JDIMENSION MCUs_per_row;
    JDIMENSION MCU_rows_in_scan;
    int blocks_in_MCU;
    int MCU_membership[D_MAX_BLOCKS_IN_MCU];
    int Ss, Se, Ah, Al;
    int unread_marker;
    struct jpeg_decomp_master* master;
    struct jpeg_d_main_controller* main;
    struct jpeg_d_coef_controller* coef;
    struct jpeg_d_post_controller* post;
    struct jpeg_input_controller* inputctl;
    struct jpeg_marker_reader* marker;
    struct jpeg_entropy_decoder* entropy;
    struct jpeg_inverse_dct* idct;
    struct jpeg_upsampler* upsample;
    struct jpeg_color_deconverter* cconvert;
    struct jpeg_color_quantizer* cquantize;
};

#define JMSG_LENGTH_MAX 200
#define JMSG_STR_PARM_MAX 80
struct jpeg_error_mgr {
    void (*error_exit)(j_common_ptr cinfo);
    void (*emit_message)(j_common_ptr cinfo, int msg_level);
    void (*output_message)(j_common_ptr cinfo);
    void (*format_message)(j_common_ptr cinfo, char* buffer);
    void (*reset_error_mgr)(j_common_ptr cinfo);
    int msg_code;
    union {
        int i[8];
        char s[JMSG_STR_PARM_MAX];
    } msg_parm;
    int trace_level;
    long num_warnings;
    const char* const* jpeg_message_table;
    int last_jpeg_message;
    const char* const* addon_message_table;
    int first_addon_message;
    int last_addon_message;
};

struct jpeg_progress_mgr {
    void (*progress_monitor)(j_common_ptr cinfo);
    long pass_counter;
    long pass_limit;
    int completed_passes;
    int total_passes;
};

struct jpeg_destination_mgr {
    JOCTET* next_output_byte;
    size_t free_in_buffer;
    void (*init_destination)(j_compress_ptr cinfo);
    boolean (*empty_output_buffer)(j_compress_ptr cinfo);
    void (*term_destination)(j_compress_ptr cinfo);
};

struct jpeg_source_mgr {
    const JOCTET* next_input_byte;
    size_t bytes_in_buffer;
    void (*init_source)(j_decompress_ptr cinfo);
    boolean (*fill_input_buffer)(j_decompress_ptr cinfo);
    void (*skip_input_data)(j_decompress_ptr cinfo, long num_bytes);
    boolean (*resync_to_restart)(j_decompress_ptr cinfo, int desired);
    void (*term_source)(j_decompress_ptr cinfo);
};

#define JPOOL_PERMANENT 0
#define JPOOL_IMAGE 1
struct jpeg_memory_mgr {
    void* (*alloc_small)(j_common_ptr cinfo, int pool_id, size_t sizeofobject);
    /* further members exist in the library; the glue only calls alloc_small through this pointer */
};

#define JPEG_SUSPENDED 0
#define JPEG_HEADER_OK 1
#define JPEG_HEADER_TABLES_ONLY 2
#define JPEG_RST0 0xD0
#define JPEG_EOI 0xD9
#define JPEG_APP0 0xE0
#define JPEG_COM 0xFE

struct jpeg_error_mgr* jpeg_std_error(struct jpeg_error_mgr* err);
void jpeg_CreateDecompress(j_decompress_ptr cinfo, int version, size_t structsize);
#define jpeg_create_decompress(cinfo) jpeg_CreateDecompress((cinfo), JPEG_LIB_VERSION, (size_t)sizeof(struct jpeg_decompress_struct))
void jpeg_destroy_decompress(j_decompress_ptr cinfo);
void jpeg_destroy(j_common_ptr cinfo);
void jpeg_abort(j_common_ptr cinfo);
int jpeg_read_header(j_decompress_ptr cinfo, boolean require_image);
boolean jpeg_start_decompress(j_decompress_ptr cinfo);
JDIMENSION jpeg_read_scanlines(j_decompress_ptr cinfo, JSAMPARRAY scanlines, JDIMENSION max_lines);
JDIMENSION jpeg_skip_scanlines(j_decompress_ptr cinfo, JDIMENSION num_lines);
void jpeg_crop_scanline(j_decompress_ptr cinfo, JDIMENSION* xoffset, JDIMENSION* width);
boolean jpeg_finish_decompress(j_decompress_ptr cinfo);
void jpeg_calc_output_dimensions(j_decompress_ptr cinfo);
boolean jpeg_resync_to_restart(j_decompress_ptr cinfo, int desired);

#ifdef __cplusplus
}
#endif
#endif
