/* oracle/shim/libusb.h -- TEST INFRASTRUCTURE.  Core/arch/linux/usb_device.h:24 includes <libusb.h> for types none of
 * the declarations it exposes to libsddc/libsddc.cpp use; the mock USB layer of oracle/ref_libsddc_driver.cpp needs no
 * libusb (none is installed here), so this header is intentionally empty. */
#ifndef ORACLE_SHIM_LIBUSB_H
#define ORACLE_SHIM_LIBUSB_H
#endif
