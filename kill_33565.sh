ssh test@127.0.0.1 "kill -9 1616"
ssh test@127.0.0.1 "kill -9 1617"
kill -9 1591
