ssh test@127.0.0.1 "kill -9 1602"
kill -9 1591
