/* Host shim for the firmware's main.h (TEST INFRASTRUCTURE, written for this repo).
 * The real firmware/src/main.h pulls in the STM32 HAL; audioprocessing.c only needs
 * the handful of names below.  Everything here is a no-op stand-in. */
#pragma once
#include <stdint.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { int unused; } UART_HandleTypeDef;
extern UART_HandleTypeDef huart1;
extern UART_HandleTypeDef huart4;

#define LED_ON()  do { } while (0)
#define LED_OFF() do { } while (0)

void Error_Handler(void);
void mainSetPrintfUart(UART_HandleTypeDef *p);
