ssh test@127.0.0.1 "kill -9 5426"
ssh test@127.0.0.1 "kill -9 5427"
kill -9 5044
