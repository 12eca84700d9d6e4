package com.fenbi.mp4j.exception;

/** The one checked exception of the API (same role as the reference's Mp4jException). */
public class Mp4jException extends Exception {
    public Mp4jException(String message) { super(message); }
    public Mp4jException(Throwable cause) { super(cause); }
    public Mp4jException(String message, Throwable cause) { super(message, cause); }
}
