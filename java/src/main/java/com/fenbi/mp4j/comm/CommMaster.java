package com.fenbi.mp4j.comm;

import com.fenbi.mp4j.exception.Mp4jException;

/** Coordination + log aggregation service; the server itself is the C++ Master of libmp4j.
 *  CLI: java com.fenbi.mp4j.comm.CommMaster &lt;slave_num&gt; &lt;port&gt; (exit code 0 iff all slaves closed 0). */
public class CommMaster {
    private final long handle;

    public CommMaster(int nSlave, int port) throws Mp4jException {
        handle = Native.masterCreate(nSlave, port, "log");
        if (handle == 0) throw new Mp4jException(Native.lastError());
    }
    public void start() throws Mp4jException {
        if (Native.masterStart(handle) != 0) throw new Mp4jException(Native.lastError());
    }
    public int getPort() { return Native.masterPort(handle); }
    /** the port slaves connect to (CommMaster.getHostPort of the reference) */
    public int getHostPort() { return getPort(); }
    /** the name slaves should use to reach this master (CommMaster.getHostName of the reference) */
    public String getHostName() {
        try {
            return java.net.InetAddress.getLocalHost().getHostName();
        } catch (java.net.UnknownHostException e) {
            return "127.0.0.1";
        }
    }
    /** Blocks until every slave closed; returns the aggregated exit code. */
    public int stop() { return Native.masterWait(handle); }

    public static void main(String[] args) throws Exception {
        if (args.length < 2) {
            System.err.println("usage: CommMaster <slave_num> <port>");
            System.exit(64);
        }
        CommMaster master = new CommMaster(Integer.parseInt(args[0]), Integer.parseInt(args[1]));
        master.start();
        int code = master.stop();
        Native.masterDestroy(master.handle);
        System.exit(code);
    }
}
