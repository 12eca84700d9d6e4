ssh test@127.0.0.1 "kill -9 8885"
kill -9 8873
