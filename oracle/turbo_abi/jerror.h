/*
 * jerror.h -- the three message codes and two macros the reference's libjpeg glue uses
 * (extensions/libjpeg_turbo/jpeg_handle.cpp:157-172).  TEST INFRASTRUCTURE ONLY.
 * The numeric values of libjpeg's message enum differ between builds, so they are looked up in the
 * library's own message table (err->jpeg_message_table) by text the first time they are needed.
 */
#ifndef B200_TURBO_ABI_JERROR_H
#define B200_TURBO_ABI_JERROR_H
#include <string.h>

#include "jpeglib.h"

static inline int b200_turbo_msg_code(j_common_ptr cinfo, const char* text)
{
    const struct jpeg_error_mgr* e = cinfo->err;
    for (int i = 0; e && e->jpeg_message_table && i <= e->last_jpeg_message; ++i)
        if (e->jpeg_message_table[i] && strcmp(e->jpeg_message_table[i], text) == 0)
            return i;
    return 0; /* JMSG_NOMESSAGE */
}
#define JERR_INPUT_EMPTY b200_turbo_msg_code((j_common_ptr)cinfo, "Empty input file")
#define JERR_FILE_READ b200_turbo_msg_code((j_common_ptr)cinfo, "Input file read error")
#define JWRN_JPEG_EOF b200_turbo_msg_code((j_common_ptr)cinfo, "Premature end of JPEG file")

#define ERREXIT(cinfo, code) ((cinfo)->err->msg_code = (code), (*(cinfo)->err->error_exit)((j_common_ptr)(cinfo)))
#define WARNMS(cinfo, code) ((cinfo)->err->msg_code = (code), (*(cinfo)->err->emit_message)((j_common_ptr)(cinfo), -1))
#endif
